// Ray sampling kernels: stratified coarse z (nerf_utils.py:49-61) and hierarchical
// inverse-CDF sampling + merge (nerf_utils.py:143-176, model_v0.py:150-156), forward and
// backward.  One warp per ray.  All arithmetic that decides an index is done with explicitly
// rounded fp32 operations in the order the oracle pins, so bins / permutations are bit-exact.
#include "common.cuh"

namespace cn {

__global__ void sample_coarse_kernel(const float *__restrict__ u, float *__restrict__ z, int64_t total, int S,
                                     double near_, double far_) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int i = (int)(idx % S);
  const double step = (far_ - near_) / (double)S;
  const float lo = (float)(near_ + (double)i * step);  // interval table built in double, cast once
  const float stepf = (float)step;
  z[idx] = __fadd_rn(lo, __fmul_rn(u[idx], stepf));  // two roundings, as in the reference graph
}

constexpr int kS = 64;   // coarse samples / importance draws per ray
constexpr int kNB = 63;  // cdf entries / bins

struct PdfShared {
  float zc[kS];
  float wc[kS];
  float cdf[kS];      // 63 used
  float keys[2 * kS]; // merged z values
  float aux[2 * kS];  // staging (sorted z / gradients)
  unsigned char prm[2 * kS];
};

// Shared forward math: fills sh.cdf, returns wsum and, for draw `uu`, the bin indices.
__device__ __forceinline__ float pdf_build_cdf(PdfShared &sh, bool &wsum_replaced) {
  float wsum = 0.f;
  for (int k = 0; k < kS - 2; ++k) wsum = __fadd_rn(wsum, __fadd_rn(sh.wc[k + 1], 1e-5f));
  wsum_replaced = (fabsf(wsum) == 0.f);
  if (wsum_replaced) wsum = 1.f;
  float c = 0.f;
  if (lane_id() == 0) sh.cdf[0] = 0.f;
  for (int k = 1; k < kNB; ++k) {
    c = __fadd_rn(c, __fdiv_rn(__fadd_rn(sh.wc[k], 1e-5f), wsum));
    if (lane_id() == 0) sh.cdf[k] = c;
  }
  __syncwarp();
  return wsum;
}

struct Draw {
  int above, below;
  float cdf_a, cdf_b, bins_a, bins_b, denom, t;
  bool denom_replaced;
};

// bins are the midpoints of the coarse z (model_v0.py:150); the op-level entry point
// cn_sample_pdf passes explicit bins instead and parks them in aux[] (kExplicitBins)
template <bool kExplicitBins = false>
__device__ __forceinline__ float bin_at(const PdfShared &sh, int k) {
  if (k >= kNB) return 0.f;
  return kExplicitBins ? sh.aux[k] : __fmul_rn(0.5f, __fadd_rn(sh.zc[k + 1], sh.zc[k]));
}

template <bool kExplicitBins = false>
__device__ __forceinline__ Draw pdf_draw(const PdfShared &sh, float uu) {
  Draw d;
  int above = 0;
  for (int k = 0; k < kNB; ++k) above += (uu >= sh.cdf[k]) ? 1 : 0;
  d.above = above;
  d.below = min(max(above - 1, 0), kNB - 1);
  d.cdf_a = above >= kNB ? 0.f : sh.cdf[above];  // out-of-range gather reads 0 (TF-GPU gather_nd)
  d.cdf_b = sh.cdf[d.below];
  d.bins_a = bin_at<kExplicitBins>(sh, above);
  d.bins_b = bin_at<kExplicitBins>(sh, d.below);
  float den = __fsub_rn(d.cdf_a, d.cdf_b);
  d.denom_replaced = den < 1e-5f;
  d.denom = d.denom_replaced ? 1.f : den;
  d.t = __fdiv_rn(__fsub_rn(uu, d.cdf_b), d.denom);
  return d;
}

__global__ void __launch_bounds__(128)
sample_pdf_fwd_kernel(const float *__restrict__ z_c, const float *__restrict__ w_c, const float *__restrict__ u,
                      int64_t n_rays, float *__restrict__ z_fine, unsigned char *__restrict__ perm,
                      unsigned char *__restrict__ above_out) {
  __shared__ PdfShared shs[4];
  const int warp = threadIdx.x >> 5;
  const int l = lane_id();
  const int64_t n = (int64_t)blockIdx.x * 4 + warp;
  if (n >= n_rays) return;
  PdfShared &sh = shs[warp];
  sh.zc[l] = z_c[n * kS + l];
  sh.zc[l + 32] = z_c[n * kS + l + 32];
  sh.wc[l] = w_c[n * kS + l];
  sh.wc[l + 32] = w_c[n * kS + l + 32];
  __syncwarp();
  bool rep;
  pdf_build_cdf(sh, rep);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = l + 32 * h;
    const float uu = u[n * kS + j];
    const Draw d = pdf_draw(sh, uu);
    const float zn = __fadd_rn(d.bins_b, __fmul_rn(d.t, __fsub_rn(d.bins_a, d.bins_b)));
    sh.keys[kS + j] = zn;
    sh.keys[j] = sh.zc[j];
    above_out[n * kS + j] = (unsigned char)d.above;
  }
  __syncwarp();
  // stable rank sort of the 128 keys (ties broken by source index == stable sort of the concat)
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const int e = l + 32 * h;
    const float key = sh.keys[e];
    int rank = 0;
    for (int f = 0; f < 2 * kS; ++f) {
      const float o = sh.keys[f];
      rank += (o < key || (o == key && f < e)) ? 1 : 0;
    }
    sh.aux[rank] = key;
    sh.prm[rank] = (unsigned char)e;
  }
  __syncwarp();
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const int e = l + 32 * h;
    z_fine[n * 2 * kS + e] = sh.aux[e];
    perm[n * 2 * kS + e] = sh.prm[e];
  }
}

// op-level form of sample_pdf (nerf_utils.py:143-176): explicit bins [n,63] and weights [n,62]
__global__ void __launch_bounds__(128)
sample_pdf_op_kernel(const float *__restrict__ bins, const float *__restrict__ weights, const float *__restrict__ u,
                     int64_t n_rays, float *__restrict__ samples, int32_t *__restrict__ above_out) {
  __shared__ PdfShared shs[4];
  const int warp = threadIdx.x >> 5;
  const int l = lane_id();
  const int64_t n = (int64_t)blockIdx.x * 4 + warp;
  if (n >= n_rays) return;
  PdfShared &sh = shs[warp];
  for (int k = l; k < kNB; k += 32) sh.aux[k] = bins[n * kNB + k];
  for (int k = l; k < kS - 2; k += 32) sh.wc[k + 1] = weights[n * (kS - 2) + k];  // wc[1..62] like w_c[1:-1]
  __syncwarp();
  bool rep;
  pdf_build_cdf(sh, rep);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = l + 32 * h;
    const Draw d = pdf_draw<true>(sh, u[n * kS + j]);
    samples[n * kS + j] = __fadd_rn(d.bins_b, __fmul_rn(d.t, __fsub_rn(d.bins_a, d.bins_b)));
    if (above_out) above_out[n * kS + j] = d.above;
  }
}

__global__ void __launch_bounds__(128)
sample_pdf_bwd_kernel(const float *__restrict__ z_c, const float *__restrict__ w_c, const float *__restrict__ u,
                      const unsigned char *__restrict__ perm, const float *__restrict__ dz_fine, int64_t n_rays,
                      float *__restrict__ dw_c) {
  __shared__ PdfShared shs[4];
  __shared__ float dcdf_s[4][kS];
  const int warp = threadIdx.x >> 5;
  const int l = lane_id();
  const int64_t n = (int64_t)blockIdx.x * 4 + warp;
  if (n >= n_rays) return;
  PdfShared &sh = shs[warp];
  float *dcdf = dcdf_s[warp];
  sh.zc[l] = z_c[n * kS + l];
  sh.zc[l + 32] = z_c[n * kS + l + 32];
  sh.wc[l] = w_c[n * kS + l];
  sh.wc[l + 32] = w_c[n * kS + l + 32];
  dcdf[l] = 0.f;
  dcdf[l + 32] = 0.f;
  // route dz_fine through the sort permutation to the importance draws
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const int s = l + 32 * h;
    const int e = perm[n * 2 * kS + s];
    if (e >= kS) sh.aux[e - kS] = dz_fine[n * 2 * kS + s];
  }
  __syncwarp();
  bool wsum_replaced;
  const float wsum = pdf_build_cdf(sh, wsum_replaced);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = l + 32 * h;
    const float uu = u[n * kS + j];
    const Draw d = pdf_draw(sh, uu);
    const float g = sh.aux[j];
    const float dt = g * (d.bins_a - d.bins_b);  // z_new = bins_b + t (bins_a - bins_b)
    const float num = uu - d.cdf_b;
    float g_b = -dt / d.denom;
    if (!d.denom_replaced) {
      const float q = dt * num / (d.denom * d.denom);
      g_b += q;
      if (d.above < kNB) atomicAdd(&dcdf[d.above], -q);
    }
    atomicAdd(&dcdf[d.below], g_b);
  }
  __syncwarp();
  // cdf_k = sum_{i<k} pdf_i  ->  dpdf_j = sum_{k>j} dcdf_k ; pdf_j = sw_j / wsum, sw_j = w_c[j+1] + 1e-5
  float dpdf[2], sw[2];
  float dot = 0.f;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = l + 32 * h;
    float acc = 0.f;
    if (j < kS - 2)
      for (int k = j + 1; k < kNB; ++k) acc += dcdf[k];
    dpdf[h] = acc;
    sw[h] = (j < kS - 2) ? sh.wc[j + 1] + 1e-5f : 0.f;
    dot += acc * sw[h];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = l + 32 * h;
    if (j < kS - 2) {
      float g = dpdf[h] / wsum;
      if (!wsum_replaced) g -= dot / (wsum * wsum);
      dw_c[n * kS + j + 1] += g;
    }
  }
}

}  // namespace cn

extern "C" int cn_sample_coarse(const float *u, float *z, int64_t n_rays, int32_t S, double near_, double far_,
                                void *stream) {
  using namespace cn;
  CN_REQUIRE(u && z, "cn_sample_coarse: null pointer");
  CN_REQUIRE(n_rays >= 0 && S > 0, "cn_sample_coarse: n_rays %lld S %d", (long long)n_rays, S);
  const int64_t total = n_rays * S;
  if (total == 0) return 0;
  const int threads = 256;
  sample_coarse_kernel<<<(unsigned)((total + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
      u, z, total, S, near_, far_);
  return check_launch("cn_sample_coarse");
}

extern "C" int cn_sample_pdf_fwd(const float *z_c, const float *w_c, const float *u, int64_t n_rays, float *z_fine,
                                 uint8_t *perm, uint8_t *above, void *stream) {
  using namespace cn;
  CN_REQUIRE(z_c && w_c && u && z_fine && perm && above, "cn_sample_pdf_fwd: null pointer");
  CN_REQUIRE(n_rays >= 0, "cn_sample_pdf_fwd: n_rays %lld", (long long)n_rays);
  if (n_rays == 0) return 0;
  sample_pdf_fwd_kernel<<<(unsigned)((n_rays + 3) / 4), 128, 0, (cudaStream_t)stream>>>(z_c, w_c, u, n_rays, z_fine,
                                                                                       perm, above);
  return check_launch("cn_sample_pdf_fwd");
}

extern "C" int cn_sample_pdf(const float *bins, const float *weights, const float *u, int64_t n_rays, float *samples,
                             int32_t *above, void *stream) {
  using namespace cn;
  CN_REQUIRE(bins && weights && u && samples, "cn_sample_pdf: null pointer");
  CN_REQUIRE(n_rays >= 0, "cn_sample_pdf: n_rays %lld", (long long)n_rays);
  if (n_rays == 0) return 0;
  sample_pdf_op_kernel<<<(unsigned)((n_rays + 3) / 4), 128, 0, (cudaStream_t)stream>>>(bins, weights, u, n_rays,
                                                                                      samples, above);
  return check_launch("cn_sample_pdf");
}

extern "C" int cn_sample_pdf_bwd(const float *z_c, const float *w_c, const float *u, const uint8_t *perm,
                                 const float *dz_fine, int64_t n_rays, float *dw_c, void *stream) {
  using namespace cn;
  CN_REQUIRE(z_c && w_c && u && perm && dz_fine && dw_c, "cn_sample_pdf_bwd: null pointer");
  CN_REQUIRE(n_rays >= 0, "cn_sample_pdf_bwd: n_rays %lld", (long long)n_rays);
  if (n_rays == 0) return 0;
  sample_pdf_bwd_kernel<<<(unsigned)((n_rays + 3) / 4), 128, 0, (cudaStream_t)stream>>>(z_c, w_c, u, perm, dz_fine,
                                                                                       n_rays, dw_c);
  return check_launch("cn_sample_pdf_bwd");
}
