// cn_adam_step: the parameter update of MVVNeRFRenderer.train_step as ONE kernel over the flat buffers --
//   optimize()            (nerf_utils.py:8-12):   per-element clip_by_value(g, -clip, clip)
//   WarmupScheduler       (nerf_utils.py:288-300): lr(t) = t / warmup * target (t <= warmup), target, 0.1 target
//   tf.keras.optimizers.Adam (train_nerf.py:23):   m, v moments, lr_t = lr sqrt(1 - b2^t) / (1 - b1^t),
//                                                  p -= lr_t m / (sqrt(v) + eps)
// The step counter lives on the DEVICE (a double), so a CUDA graph that recorded this launch replays the schedule and
// the bias corrections of every later step: the update reads t = counter + 1, a second one-thread launch stores t.
#include "common.cuh"

namespace cn {

__global__ void adam_step_kernel(float *__restrict__ params, float *__restrict__ grad, float *__restrict__ m,
                                 float *__restrict__ v, int64_t n_rows, int64_t row_stride, int64_t n_update,
                                 const double *__restrict__ counter, double target_lr, double warmup,
                                 double scale_down_after, float grad_scale, float clip, double beta1, double beta2,
                                 float eps) {
  __shared__ float s_lr_t;
  if (threadIdx.x == 0) {
    const double t = *counter + 1.0;
    const double lr = t <= warmup ? t / warmup * target_lr : (t <= scale_down_after ? target_lr : 0.1 * target_lr);
    s_lr_t = (float)(lr * sqrt(1.0 - pow(beta2, t)) / (1.0 - pow(beta1, t)));
  }
  __syncthreads();
  const float lr_t = s_lr_t;
  const float b1 = (float)beta1, b2 = (float)beta2, c1 = (float)(1.0 - beta1), c2 = (float)(1.0 - beta2);
  const int64_t total = n_rows * n_update;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t idx = (i / n_update) * row_stride + i % n_update;
    float g = grad[idx] * grad_scale;
    if (clip > 0.f) g = fminf(fmaxf(g, -clip), clip);
    grad[idx] = g;                                   // the clipped gradient stays visible to the caller
    const float mi = fmaf(c1, g, b1 * m[idx]);
    const float vi = fmaf(c2, g * g, b2 * v[idx]);
    m[idx] = mi;
    v[idx] = vi;
    params[idx] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

__global__ void counter_inc_kernel(double *counter) { *counter += 1.0; }

}  // namespace cn

extern "C" int cn_adam_step(float *params, float *grad, float *m, float *v, int64_t n_rows, int64_t row_stride,
                            int64_t n_update, double *step_counter, double target_lr, double warmup_steps,
                            double scale_down_after, float grad_scale, float clip, double beta1, double beta2,
                            double eps, void *stream) {
  using namespace cn;
  CN_REQUIRE(params && grad && m && v && step_counter, "cn_adam_step: null pointer");
  CN_REQUIRE(n_rows >= 0 && n_update >= 0 && row_stride >= n_update, "cn_adam_step: n_rows=%lld row_stride=%lld n_update=%lld",
             (long long)n_rows, (long long)row_stride, (long long)n_update);
  CN_REQUIRE(warmup_steps >= 1.0 && beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0,
             "cn_adam_step: warmup_steps=%g beta1=%g beta2=%g", warmup_steps, beta1, beta2);
  const int64_t total = n_rows * n_update;
  if (total > 0) {
    const int64_t want = (total + 255) / 256;
    const unsigned grid = (unsigned)(want < 8 * num_sms() ? want : 8 * num_sms());
    adam_step_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(params, grad, m, v, n_rows, row_stride, n_update,
                                                             step_counter, target_lr, warmup_steps, scale_down_after,
                                                             grad_scale, clip, beta1, beta2, (float)eps);
  }
  counter_inc_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_counter);
  return check_launch("cn_adam_step");
}
