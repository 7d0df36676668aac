"""Frame-level data parallelism for the camera/time batch.

The reference trains on `num_cams_per_iter` (camera, timestamp) frames per iteration in a serial Python
loop on one GPU (train.py:224-247); frames are independent given the parameters, so they shard over
ranks with ONE exchange step: the sum of the parameter gradients.  Parameters stay replicated.

`FlatGrads` makes every parameter's `.grad` a view into one flat fp32 buffer, so that exchange is a single
all-reduce with no packing kernel (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
import torch


def shard_frames(num_frames, world_size, rank):
    """Frame k of the time-sorted batch goes to rank k mod G (train.py:148-154 sorts the batch in time)."""
    return list(range(rank, num_frames, world_size))


def frame_loss_weights(num_frames, lambda_dssim):
    """Per-frame (L1 weight, SSIM weight) reproducing the reference's cumulative-Ll1 quirk
    (train.py:245-250: `Ll1` accumulates across frames before entering the loss, then loss /= k)."""
    T = num_frames
    return [((T - j) * (1.0 - lambda_dssim) / T, lambda_dssim / T) for j in range(T)]


class GraphedStep:
    """Capture one whole step of the path — DeformModel.step -> render -> loss -> backward, written against STATIC
    tensors (parameters, camera tensors, ground truth, the flat gradient buffer) — into a CUDA graph and replay it.
    A frame of this path is ~45 kernel launches issued from Python through five autograd Functions (~1.35 ms of
    host time per frame, measured with a 2k-Gaussian frame) for ~1.6 ms of GPU work at 100k Gaussians, so an eager
    loop is about to become launch-bound; a replay costs one launch.

        step = GraphedStep(lambda: train_frame())      # runs fn a few times eagerly, then captures it
        for it in ...:  upload_inputs(); out = step()  # out = whatever fn returned at capture (static tensors)

    Requirements on fn: no host synchronisation (size the rasterizer's binning workspace with a
    rasterizer.PAIR_CAPACITY_POLICY that has seen the shapes, e.g. GrowingPairCapacity after the warm-up calls), inputs
    updated in place.  The deformation network re-packs its weights inside the graph, so parameter updates made
    between replays (optimizer steps) are seen.  Status words / pair-count headers copied to pinned memory inside the
    graph are examined by check() (call it where the host synchronises anyway)."""

    def __init__(self, fn, warmup=3, checks=()):
        self.fn, self.checks = fn, list(checks)
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()

    def __call__(self):
        self.graph.replay()
        return self.out

    def check(self):
        for c in self.checks:
            c()


def capturing():
    return torch.cuda.is_available() and torch.cuda.is_current_stream_capturing()


class FlatGrads:
    """`params` in exchange order.  With `early` = the number of leading parameters whose gradients are
    complete early in backward (the Gaussian parameters: autograd accumulates them right after the rasterizer
    backward, before the deformation-MLP backward runs), `all_reduce()` becomes two collectives: the leading
    slice is reduced asynchronously as soon as its last gradient has been accumulated — overlapping the MLP
    backward kernels — and only the small trailing slice (the MLP gradients) is reduced at the end."""

    def __init__(self, params, early=0):
        self.params = [p for p in params]
        dev = self.params[0].device
        # every view starts on a 16-byte boundary so the kernels' vector reductions (red.global.add.v4.f32)
        # apply; the padding floats stay zero and ride along in the all-reduce
        sizes = [(p.numel() + 3) // 4 * 4 for p in self.params]
        self.flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        o = 0
        self.early_len = 0
        for i, (p, n) in enumerate(zip(self.params, sizes)):
            p.grad = self.flat[o:o + p.numel()].view_as(p)
            o += n
            if i + 1 == early:
                self.early_len = o
        self._early_n, self._seen, self._work, self._group = int(early), 0, None, None
        self.frames_per_reduce = 1  # local frames whose backward passes accumulate before one exchange
        # renderer.FUSED_GRAD_ACCUMULATION adds the Gaussian gradients inside the kernels (no AccumulateGrad, so no
        # hook): render() then reports every finished frame through on_frame_grads(); frames rendered per exchange
        self.frames_per_exchange = 1
        self._frames_seen = 0
        if early:
            for p in self.params[:early]:
                p.register_post_accumulate_grad_hook(self._on_grad)

    def _on_grad(self, _p):
        self._seen += 1
        if self._seen == self._early_n * self.frames_per_reduce:
            self._seen = 0
            self._reduce_early()

    def on_frame_grads(self):
        """renderer.ON_GAUSSIAN_GRADS callback: one frame's Gaussian-parameter gradients are complete in `flat`."""
        if not self._early_n:
            return
        self._frames_seen += 1
        if self._frames_seen == self.frames_per_exchange:
            self._frames_seen = 0
            self._reduce_early()

    def _reduce_early(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self._group) > 1:
            self._work = dist.all_reduce(self.flat[:self.early_len], op=dist.ReduceOp.SUM, group=self._group,
                                         async_op=True)

    def zero(self):
        self.flat.zero_()

    def all_reduce(self, group=None):
        """Sum the gradients over ranks (call after backward of this rank's last frame)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            if self._work is not None:  # leading slice already in flight since mid-backward
                dist.all_reduce(self.flat[self.early_len:], op=dist.ReduceOp.SUM, group=group)
                self._work.wait()
                self._work = None
            else:
                dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        return self.flat


def add_densification_stats(gaussians, render_pkg, update_max_radii=True):
    """What the training loop does with one frame's render_pkg after backward — train.py:238-239 (`max_radii2D`) and
    :283-284 -> GaussianModel.add_densification_stats (scene/gaussian_model.py:398-401) on
    `viewspace_points_densify.grad` — as ONE kernel over the Gaussians instead of boolean-mask gathers and scatters.
    `gaussians` needs the reference's `xyz_gradient_accum` [N,1], `denom` [N,1] and `max_radii2D` [N] tensors."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    radii = render_pkg["radii"]
    vs = render_pkg["viewspace_points_densify"].grad
    if vs is not None and (vs.dtype != torch.float32 or not vs.is_contiguous()):
        vs = vs.float().contiguous()
    N = int(radii.shape[0])
    acc, den, mr = gaussians.xyz_gradient_accum, gaussians.denom, gaussians.max_radii2D
    for t in (acc, den, mr):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.device != radii.device:
            raise RuntimeError("add_densification_stats: accumulators must be contiguous fp32 tensors on the "
                               "rasterizer's device")
    with torch.cuda.device(radii.device):
        _lib.check(lib.d3g_densification_stats(N, _lib.ptr(vs), _lib.ptr(radii), _lib.ptr(acc), _lib.ptr(den),
                                               _lib.ptr(mr) if update_max_radii else None,
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream)))


def reduce_densification_stats(grad_norm_accum, denom, max_radii2D, group=None):
    """Densification statistics are per-frame quantities (scene/gaussian_model.py:398-401 takes the norm of
    each frame's viewspace gradient BEFORE accumulating), so ranks reduce their local accumulators:
    SUM for the gradient-norm accumulator and the visibility count, MAX for max_radii2D (train.py:238-239)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(grad_norm_accum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)
    return grad_norm_accum, denom, max_radii2D
