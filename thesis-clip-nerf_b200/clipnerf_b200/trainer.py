"""Training step of the NeRF MLPs on the sm_100a render path: the replacement of
``MVVNeRFRenderer.train_step`` + ``optimize`` (reference ``src/lib/mvnerf/model_v0.py:186-197``,
``src/lib/mvnerf/nerf_utils.py:8-12``) for the part of the model that lives on the hot path.

One process per GPU.  All trainable NeRF variables of both nets live in ONE flat fp32 buffer
``[2, NET_PARAMS]`` whose gradient is a single flat buffer as well, so data parallelism over ray
shards is exactly one NCCL all-reduce per step (no collective is needed for rendering)."""
from __future__ import annotations

import queue
import threading
from typing import Optional

import torch
import torch.distributed as dist

from . import params as P
from ._abi import EMB_PARAMS, NET_PARAMS
from . import render as _render
from .render import render_rays


def warmup_lr(step: int, target_lr: float, warmup_steps: float = 10000.0, scale_down_after: float = 450000.0) -> float:
    """``WarmupScheduler`` (nerf_utils.py:288-300) with the arguments of train_nerf.py:22."""
    warmup_steps = max(1.0, float(warmup_steps))
    if step <= warmup_steps:
        return step / warmup_steps * target_lr
    return target_lr if step <= scale_down_after else 0.1 * target_lr


def allreduce_mean(grads, like, group=None):
    """mean over the ranks of a list of gradient tensors (``None`` entries count as zeros of the matching tensor in
    ``like``) with ONE collective on a flat buffer -- the data-parallel step of the pose-query trainers
    (``DeltaNGF.train_step``: the 582 337 GraspReadout parameters, 2.33 MB).  No-op on one rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return list(grads)
    flat = torch.cat([(torch.zeros_like(p) if g is None else g).reshape(-1) for g, p in zip(grads, like)])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    flat /= dist.get_world_size(group)
    out, o = [], 0
    for p in like:
        out.append(flat[o:o + p.numel()].view_as(p))
        o += p.numel()
    return out


class RenderTrainer:
    """Flat-buffer trainer of the two NeRF nets.

    Which variables are updated.  The reference's ``MultiOptimizer`` (``src/train_nerf.py:27-33``) holds
    ``coarse_embedding`` and ``fine_embedding`` (plus the image encoders); ``coarse_readout`` / ``fine_readout`` are in
    no optimizer group, so ``apply_gradients`` never touches them.  ``train_readouts=False`` (default) reproduces that:
    Adam runs over the first ``EMB_PARAMS`` entries of every net, the 516 readout parameters keep their values (their
    gradient is still computed, as in the reference).  ``train_readouts=True`` updates all ``NET_PARAMS``.

    ``params``: an existing ``[2, NET_PARAMS]`` leaf tensor to train in place (``MVVNeRFRenderer`` passes its own
    parameter buffer); default: a fresh glorot initialisation."""

    def __init__(self, device, seed: int = 0, target_lr: float = 1e-4, warmup_steps: float = 10000.0,
                 clip: float = 1.0, bias_std: float = 0.0, process_group=None, params: Optional[torch.Tensor] = None,
                 train_readouts: bool = False, scale_down_after: float = 450000.0):
        if params is None:
            g = torch.Generator().manual_seed(seed)
            flat = torch.stack([P.glorot_init(g, bias_std), P.glorot_init(g, bias_std)])
            self.params = flat.to(device).requires_grad_(True)          # [2, NET_PARAMS]: coarse, fine
        else:
            assert tuple(params.shape) == (2, NET_PARAMS) and params.is_contiguous() and params.dtype == torch.float32
            self.params = params
        self.m = torch.zeros_like(self.params, requires_grad=False)
        self.v = torch.zeros_like(self.params, requires_grad=False)
        self.iterations = 0
        self.target_lr, self.warmup_steps, self.clip = target_lr, max(1.0, float(warmup_steps)), clip
        self.scale_down_after = float(scale_down_after)
        self.beta1, self.beta2, self.eps = 0.9, 0.999, 1e-7           # tf.keras.optimizers.Adam defaults
        self.n_update = NET_PARAMS if train_readouts else EMB_PARAMS
        self.group = process_group
        # True: start the all-reduce inside the backward, under the feature-map sweep (render.GRAD_SYNC).  Measured on
        # 8 B200s it is no faster than the plain all-reduce behind the backward (7.40 vs 7.32 ms / step; 7.31 vs 7.33 on 2):
        # the 1.98 MB collective costs ~50 us over NVSwitch and its kernel competes with the sweep for SMs -- off by default
        # > 1 rank: the all-reduce starts inside the backward, right after the last weight-gradient kernel, and runs on
        # NCCL's stream under the feature-map sweep (8 GPUs, graph replay: 7.08 -> 7.04 ms per step, the whole cost of the collective)
        self.overlap_allreduce = True
        self._t_dev = None          # device-side step counter (float64) of apply_gradients_on_device
        self._synced = False
        self.world = dist.get_world_size(process_group) if (dist.is_available() and dist.is_initialized()) else 1

    @property
    def coarse(self):
        return self.params[0]

    @property
    def fine(self):
        return self.params[1]

    def loss_and_backward(self, inputs, features, target, u_strat, u_pdf, near, far, max_rays: Optional[int] = None):
        """forward + backward of MSE(coarse) + MSE(fine) (model_v0.py:190-194); returns (loss, outputs).
        ``max_rays``: batches with more rays per scene are rendered in ray chunks of that size whose gradients
        accumulate (rays are independent; the saved activations of one chunk are ~1.2 MB per ray at 3 views)."""
        self.params.grad = None
        self._synced = False
        nr = inputs[0].shape[1]
        if max_rays is None or nr <= max_rays:
            out = render_rays(self.params[0], self.params[1], inputs, features, u_strat, u_pdf, near, far)
            loss = torch.mean((out[0] - target) ** 2) + torch.mean((out[2] - target) ** 2)
            # one backward = the whole gradient: the all-reduce is issued inside it, under the feature-map sweep
            _render.GRAD_SYNC = self._sync_inside_backward if (self.world > 1 and self.overlap_allreduce) else None
            try:
                loss.backward()
            finally:
                _render.GRAD_SYNC = None
            return loss.detach(), out
        ray_o, ray_d, images, K, Einv = inputs
        total, outs = None, []
        for s in range(0, nr, max_rays):
            sl = slice(s, min(nr, s + max_rays))
            out = render_rays(self.params[0], self.params[1], (ray_o[:, sl], ray_d[:, sl], images, K, Einv), features,
                              u_strat[:, sl], u_pdf[:, sl], near, far)
            tgt = target[:, sl]
            loss = (((out[0] - tgt) ** 2).sum() + ((out[2] - tgt) ** 2).sum()) / target.numel()
            loss.backward()
            total = loss.detach() if total is None else total + loss.detach()
            outs.append([o.detach() for o in out])
        return total, tuple(torch.cat([o[i] for o in outs], dim=1) for i in range(4))

    def _sync_inside_backward(self, g_both):
        """render.GRAD_SYNC hook: sum of the [2, NET_PARAMS] gradient buffer over ranks, asynchronous to the
        compute stream (NCCL's own stream) so that it overlaps the kernels launched behind it"""
        self._synced = True
        return dist.all_reduce(g_both, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def allreduce_gradients(self):
        """the one collective of the path: mean of the flat gradient buffer over ray shards (already summed inside
        the backward when ``overlap_allreduce`` applied)."""
        if self.world > 1:
            if not self._synced:
                dist.all_reduce(self.params.grad, op=dist.ReduceOp.SUM, group=self.group)
            self.params.grad.div_(self.world)

    @torch.no_grad()
    def apply_gradients(self):
        """per-element clip to [-clip, clip] (nerf_utils.py:9-11) then Keras-style Adam, with torch ops and HOST
        scalars: the device-independent statement of the update (CPU / gloo tests, and the reference the one-kernel
        form ``apply_gradients_on_device`` is tested against).  Keeps the device-side counter in step."""
        g = self.params.grad
        if self.clip > 0:
            g.clamp_(-self.clip, self.clip)
        self.iterations += 1
        t = self.iterations
        lr = warmup_lr(t, self.target_lr, self.warmup_steps, self.scale_down_after)
        lr_t = lr * (1.0 - self.beta2 ** t) ** 0.5 / (1.0 - self.beta1 ** t)
        n = self.n_update
        p, m, v, g = self.params[:, :n], self.m[:, :n], self.v[:, :n], g[:, :n]
        m.mul_(self.beta1).add_(g, alpha=1.0 - self.beta1)
        v.mul_(self.beta2).addcmul_(g, g, value=1.0 - self.beta2)
        p.sub_(float(torch.tensor(lr_t, dtype=torch.float32)) * m / v.sqrt().add_(self.eps))
        if self._t_dev is not None:
            self._t_dev.fill_(float(self.iterations))

    def _counter(self):
        if self._t_dev is None:
            self._t_dev = torch.zeros((), dtype=torch.float64, device=self.params.device)
            self._t_dev.fill_(float(self.iterations))
        return self._t_dev

    @torch.no_grad()
    def apply_gradients_on_device(self):
        """the same update as ONE kernel (``cn_adam_step``) with the step counter, the warm-up schedule and the bias
        corrections evaluated on the device: what the training step launches, eagerly and recorded into a CUDA graph.
        The caller keeps ``self.iterations`` (host mirror of the counter) in step."""
        from . import ops
        ops.adam_step(self.params.data, self.params.grad, self.m, self.v, self.n_update, self._counter(),
                      self.target_lr, self.warmup_steps, self.scale_down_after, 1.0, self.clip, self.beta1, self.beta2,
                      self.eps)

    def capture(self, inputs, features, target, u_strat, u_pdf, near, far, pool=None, warmup: int = 0):
        """One ``train_step`` on FIXED device buffers recorded into a CUDA graph: ~45 kernel launches of this library,
        the loss, the gradient all-reduce (more than one rank) and the optimizer update -- replayed with a single launch, so the host's
        per-step work no longer bounds the loop.  Returns ``replay() -> loss`` (a device scalar overwritten by the next
        replay).  ``features.grad`` is written by every replay (memory owned by the graph).  The ``warmup`` eager
        steps DO update the parameters."""
        features.grad = None
        self._counter().fill_(float(self.iterations))  # (outside the graph) continue from the eager steps so far

        def step():
            features.grad = None
            loss, _ = self.loss_and_backward(inputs, features, target, u_strat, u_pdf, near, far)
            self.allreduce_gradients()         # > 1 rank: the NCCL all-reduce is recorded into the graph with the kernels
            self.apply_gradients_on_device()
            return loss
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                step()
                self.iterations += 1
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        # thread-local capture mode: the batch-producer thread of fit() pins host memory and waits for copy events
        # while this thread records (under the default global mode any such call invalidates the capture)
        with torch.cuda.graph(graph, pool=pool, capture_error_mode='thread_local'):
            loss = step()
        d_features = features.grad             # this graph's map gradient (graph-owned memory)
        t_dev = self._t_dev

        def replay():
            # ONE counter: the device-side step count is refilled from the host's before every replay (a 5 us fill on
            # the stream, no synchronisation), so eager steps taken between replays cannot leave it stale
            t_dev.fill_(float(self.iterations))
            graph.replay()
            self.iterations += 1
            features.grad = d_features         # (another graph of the same pool may have been replayed in between)
            return loss
        replay.graph = graph
        return replay

    def train_step(self, inputs, features, target, u_strat, u_pdf, near, far, max_rays: Optional[int] = None):
        loss, _ = self.loss_and_backward(inputs, features, target, u_strat, u_pdf, near, far, max_rays)
        self.allreduce_gradients()
        if self.params.is_cuda:
            self._counter().fill_(float(self.iterations))
            self.apply_gradients_on_device()       # the same kernel the graph replays: one update expression
            self.iterations += 1
        else:
            self.apply_gradients()
        return loss


class HostBatch(list):
    """a host batch (list of pinned tensors) whose buffers are recycled by its producer: ``uploaded`` is set by the
    ``DevicePrefetcher`` once the host->device copies are enqueued, ``copied`` is the CUDA event behind them"""

    def __init__(self, tensors):
        super().__init__(tensors)
        self.uploaded = threading.Event()
        self.copied = None


# One long-lived producer thread per process: starting a thread cost every fit() call ~3 ms before its second batch
# could begin (tools/diag_fit.py); jobs that arrive while it is busy get a thread of their own.
_producer_jobs = queue.Queue()
_producer_state = {'thread': None, 'busy': False}
_producer_lock = threading.Lock()


def _producer_loop():
    while True:
        job = _producer_jobs.get()
        try:
            job()
        finally:
            with _producer_lock:
                _producer_state['busy'] = False


def _producer_submit(job):
    with _producer_lock:
        if _producer_state['thread'] is None or not _producer_state['thread'].is_alive():
            _producer_state['thread'] = threading.Thread(target=_producer_loop, daemon=True)
            _producer_state['thread'].start()
            _producer_state['busy'] = False
        if _producer_state['busy']:
            threading.Thread(target=job, daemon=True).start()
            return
        _producer_state['busy'] = True
    _producer_jobs.put(job)


class PinnedBatchQueue:
    """``sequence[i]`` for ``i in indices`` -> pinned host batches, produced by ONE background thread a few batches
    ahead of the consumer -- the role of Keras' ``OrderedEnqueuer`` (workers = 1) under ``fit``
    (``src/train_nerf.py:52`` -> ``MVNeRFDataGenerator.__getitem__``).  One thread keeps the order of the numpy global
    RNG draws of the generator (``bbox_biased_sample``, ``np.random.choice``) deterministic.  ``flatten(item)`` maps
    the generator's item to a flat list of arrays; a ring of pinned buffer sets is recycled (a set is rewritten only
    after its upload has completed)."""

    def __init__(self, sequence, indices, flatten, depth: int = 3, getter=None, ring=None):
        self.sequence, self.indices, self.flatten = sequence, list(indices), flatten
        self.getter = getter if getter is not None else sequence.__getitem__
        self.q = queue.Queue(maxsize=depth)
        # pinned buffer sets; a caller that runs many epochs passes its own list so that they are allocated once
        self.ring = ring if ring is not None and len(ring) == depth + 3 else [None] * (depth + 3)
        if ring is not None and ring is not self.ring:
            ring[:] = self.ring
            self.ring = ring
        self.error = None
        # The FIRST batch is assembled by the calling thread, the rest by the background thread (same order of the
        # numpy RNG draws): the caller's core is awake, a freshly started thread on an idle core needed 4 ms instead
        # of 1.1 ms for it (tools/diag_fit.py) -- and the device sits idle until the first batch exists
        self.first = 0
        if self.indices:
            try:
                self._produce(0, self.indices[0])
                self.first = 1
            except BaseException as e:
                self.error = e
                self.first = len(self.indices)
        _producer_submit(self._run)

    def _produce(self, n, i):
        arrays = [torch.as_tensor(a) for a in self.flatten(self.getter(i))]
        arrays = [a.to(torch.float32) if a.dtype == torch.float64 else a for a in arrays]
        k = n % len(self.ring)
        old = self.ring[k]
        if old is not None:
            old.uploaded.wait()
            if old.copied is not None:
                old.copied.synchronize()
        if old is None or any(o.shape != a.shape or o.dtype != a.dtype for o, a in zip(old, arrays)):
            bufs = [torch.empty(a.shape, dtype=a.dtype).pin_memory() for a in arrays]
        else:
            bufs = list(old)
        for b_, a in zip(bufs, arrays):
            b_.copy_(a)
        self.ring[k] = HostBatch(bufs)
        self.q.put(self.ring[k])

    def _run(self):
        try:
            for n, i in enumerate(self.indices):
                if n >= self.first:
                    self._produce(n, i)
        except BaseException as e:          # surfaced in the consumer
            self.error = e
        self.q.put(None)

    def __iter__(self):
        while True:
            item = self.q.get()
            if item is None:
                if self.error is not None:
                    raise self.error
                return
            yield item


class DevicePrefetcher:
    """Host batches (pinned tensors) -> device batches, one batch ahead, on a dedicated copy stream.

    The role Keras' ``fit`` gives its ``Sequence`` enqueuer thread in the reference
    (``src/train_nerf.py:132`` -> ``MVNeRFDataGenerator.__getitem__``): the host->device copy of
    batch i+1 runs on the copy engine while the kernels of batch i run.  Every batch is copied exactly
    once; two device buffer sets are recycled, guarded by events (a set is overwritten only after the
    step that read it has finished)."""

    def __init__(self, batches, device, depth: int = 2, slots=None, copy_stream=None):
        self.it = iter(batches)
        self.device = device
        self.copy_stream = copy_stream if copy_stream is not None else torch.cuda.Stream(device)
        # device tensors of each buffer set; a caller that needs stable addresses (CUDA graphs) passes its own list
        self.slots = slots if slots is not None else [None] * depth
        depth = len(self.slots)
        self.free = [None] * depth             # event: the consumer is done with the set
        self.k = 0
        self.next = None
        self.exhausted = False

    def _upload(self):
        try:
            host = next(self.it)
        except StopIteration:
            self.next = None
            self.exhausted = True
            return
        k = self.k % len(self.slots)
        self.k += 1
        with torch.cuda.stream(self.copy_stream):
            if self.free[k] is not None:
                self.copy_stream.wait_event(self.free[k])
            if self.slots[k] is None or any(d.shape != h.shape for d, h in zip(self.slots[k], host)):
                self.slots[k] = [torch.empty_like(t, device=self.device) for t in host]
            for d, h in zip(self.slots[k], host):
                d.copy_(h, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(self.copy_stream)
        if isinstance(host, HostBatch):
            host.copied = ready
            host.uploaded.set()
        self.next = (k, ready)

    def __iter__(self):
        return self

    def __next__(self):
        # The upload of batch i is enqueued here, i.e. right after the host has enqueued step i-1: it runs on the copy
        # stream while step i-1 runs on the device (the host is ahead of the device in steady state), and the FIRST
        # step of a loop is enqueued as soon as its own batch exists -- not after the producer has made a second one
        # (that cost every fit() call ~2 ms of an idle device per 1.2 ms batch the producer needs)
        if self.next is None and not self.exhausted:
            self._upload()
        if self.next is None:
            raise StopIteration
        k, ready = self.next
        self.next = None
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ready)
        if self.free[k] is None:
            self.free[k] = torch.cuda.Event()
        return k, self.slots[k], self.free[k]


class _LossReader:
    """losses of the steps, copied to pinned host memory asynchronously and read two steps late (the launch queue
    holds up to two steps, so host jitter of a whole step does not idle the device)"""

    def __init__(self, device, ring=None):
        self.cur = torch.cuda.current_stream(device)
        # pinned once: no host allocation in the loop -- and, handed in by a caller that loops over many fit() calls,
        # none per call either (cudaHostAlloc + stream creation cost the first launch of every call ~3 ms)
        self.ring = ring if ring is not None else torch.empty(4, dtype=torch.float32, pin_memory=True)
        self.k, self.pending, self.losses = 0, [], []

    def push(self, loss):
        host_loss = self.ring[self.k % 4]
        self.k += 1
        host_loss.copy_(loss, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(self.cur)
        self.pending.append((host_loss, ev))
        if len(self.pending) > 2:
            h, e = self.pending.pop(0)
            e.synchronize()
            self.losses.append(float(h))

    def finish(self):
        for h, e in self.pending:
            e.synchronize()
            self.losses.append(float(h))
        self.pending = []
        return self.losses


def _step_operands(trainer, k, batch, features, draws, state):
    """(inputs5, features tensor, target, u_strat, u_pdf) of device buffer set ``k``.  ``features``: a tensor, or a
    callable ``images [b,nv,H,W,3] -> [b,nv,H,W,256]`` (the upstream encoders; its result is copied into a per-set
    buffer when it does not keep its address, so that a recorded graph finds it).  ``draws``: None (the batch carries
    u_strat / u_pdf) or a torch.Generator the stratified / inverse-CDF draws are taken from on the device."""
    inputs, target = list(batch[:5]), batch[5]
    if inputs[2].dtype == torch.uint8:
        # source images uploaded as the dataset's uint8 pixels (MVNeRFDataGenerator.get_data_compact): the reference's
        # `colors / 255.0` runs on the device, into a float buffer of this set (stable address for a recorded graph)
        from . import ops
        img = state.get(('img', k))
        shape = tuple(inputs[2].shape[:-1]) + (3,)
        if img is None or tuple(img.shape) != shape:
            img = state[('img', k)] = torch.empty(shape, dtype=torch.float32, device=inputs[2].device)
        inputs[2] = ops.images_u8_to_f32(inputs[2], out=img)
    if draws is None:
        u_s, u_p = batch[6], batch[7]
    else:
        us = state.setdefault(('u', k), None)
        shape = tuple(inputs[0].shape[:2]) + (64,)
        if us is None or tuple(us[0].shape) != shape:
            us = state[('u', k)] = [torch.empty(shape, device=inputs[0].device) for _ in range(2)]
        for u in us:
            torch.rand(shape, generator=draws, out=u)
        u_s, u_p = us
    if callable(features):
        f = features(inputs[2])
        buf = state.get(('f', k))
        if buf is None or buf.shape != f.shape:
            buf = state[('f', k)] = f.detach()          # shares f's storage: an encoder that keeps its output buffer costs no copy
            buf.requires_grad_(f.requires_grad)
        elif f.data_ptr() != buf.data_ptr():
            with torch.no_grad():
                buf.copy_(f)
        state[('f_out', k)] = f
        feats = buf
    else:
        feats = features
    return inputs, feats, target, u_s, u_p


def fit_host_batches_graph(trainer: RenderTrainer, batches, features, near, far, draws=None, after_step=None):
    """``fit_host_batches`` with the training step replayed from CUDA graphs (one per device buffer set of the
    prefetcher, sharing one memory pool): per step the host enqueues the uploads, one graph launch and the loss
    read-back.  Fixed batch shapes; on more than one rank the NCCL all-reduce is part of the graphs (drop
    ``trainer._graphs`` before destroying the process group).  ``features`` / ``draws``: see ``_step_operands``;
    ``after_step(k, features_tensor, encoder_output)`` runs after every step (e.g. the encoders' own backward from
    ``features_tensor.grad``)."""
    dev = trainer.params.device
    cur = torch.cuda.current_stream(dev)
    graphs = trainer.__dict__.setdefault('_graphs', {})
    slots = trainer.__dict__.setdefault('_graph_slots', [None, None])
    state = trainer.__dict__.setdefault('_graph_state', {})
    if 'loss_ring' not in state:           # per-trainer, not per call: pinned loss ring and the copy stream
        state['loss_ring'] = torch.empty(4, dtype=torch.float32, pin_memory=True)
        state['copy_stream'] = torch.cuda.Stream(dev)
    reader = _LossReader(dev, state['loss_ring'])
    for k, batch, done in DevicePrefetcher(batches, dev, slots=slots, copy_stream=state['copy_stream']):
        inputs, feats, target, u_s, u_p = _step_operands(trainer, k, batch, features, draws, state)
        key = tuple(t.data_ptr() for t in batch) + (feats.data_ptr(), u_s.data_ptr(), u_p.data_ptr())
        if trainer.iterations == 0:
            # the first step of a trainer's life runs eagerly (lazy initialisations must not happen under capture)
            loss = trainer.train_step(inputs, feats, target, u_s, u_p, near, far)
        else:
            if key not in graphs:
                pool = next(iter(graphs.values())).graph.pool() if graphs else None
                torch.cuda.synchronize(dev)
                graphs[key] = trainer.capture(inputs, feats, target, u_s, u_p, near, far, pool=pool)
                torch.cuda.synchronize(dev)
            loss = graphs[key]()
        if after_step is not None:
            after_step(k, feats, state.get(('f_out', k)))
        done.record(cur)
        reader.push(loss)
    return reader.finish()


def fit_host_batches(trainer: RenderTrainer, batches, features, near, far, max_rays: Optional[int] = None, draws=None,
                     after_step=None):
    """Training loop over HOST batches ``(ray_o, ray_d, images, K, E^-1, target[, u_strat, u_pdf])`` in pinned
    memory: prefetched uploads, one eager ``train_step`` per batch, the loss of every step copied back to the host
    asynchronously.  Returns the losses."""
    dev = trainer.params.device
    cur = torch.cuda.current_stream(dev)
    reader = _LossReader(dev)
    state = {}
    for k, batch, done in DevicePrefetcher(batches, dev):
        inputs, feats, target, u_s, u_p = _step_operands(trainer, k, batch, features, draws, state)
        feats.grad = None
        loss = trainer.train_step(inputs, feats, target, u_s, u_p, near, far, max_rays)
        if after_step is not None:
            after_step(k, feats, state.get(('f_out', k)))
        done.record(cur)
        reader.push(loss)
    return reader.finish()
