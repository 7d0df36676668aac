"""Drop-in for the reference's deformation networks and their wrappers.

`DeformNetwork` / `DeformNetworkBaseline` mirror utils/time_utils.py:56-201 (same constructor
arguments, same parameter names / shapes / construction order, so `state_dict()` files and seeded
initialisation are interchangeable); `DeformModel` / `DeformModelBaseline` mirror
scene/deform_model.py:253-337 (`.deform`, `.optimizer`, `.spatial_lr_scale`, `.step`,
`.train_setting`, `.save_weights`, `.load_weights`, `.update_learning_rate`).

forward() does no torch math: positional encoding, the 8x256 trunk, the heads and the se(3) map run in
the fused tcgen05 kernels of libd3g.so through the C ABI (d3g_mlp_forward / d3g_mlp_backward).
"""
import ctypes as C
import os

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .optim import Adam as _FusedAdam  # torch.optim.Adam subclass: same state/param_groups, one fused step kernel

HEAD_WARP, HEAD_ROT, HEAD_SCALE = 1, 2, 4


def _param_list(net):
    """(name, tensor) in the fixed order of the C structs d3g_mlp_params / d3g_mlp_grads.  Cached on the module
    (the Parameter objects survive .cuda() / load_state_dict(); _apply() drops the cache anyway)."""
    cached = net.__dict__.get("_named_cache")
    if cached is not None:
        return cached
    out = _param_list_uncached(net)
    net.__dict__["_named_cache"] = out
    return out


def _param_list_uncached(net):
    out = []
    if net.is_blender:
        out += [("timenet0_w", net.timenet[0].weight), ("timenet0_b", net.timenet[0].bias),
                ("timenet2_w", net.timenet[2].weight), ("timenet2_b", net.timenet[2].bias)]
    for i in range(8):
        out += [("linear_w%d" % i, net.linear[i].weight), ("linear_b%d" % i, net.linear[i].bias)]
    if net.is_6dof:
        out += [("branch_w_w", net.branch_w.weight), ("branch_w_b", net.branch_w.bias),
                ("branch_v_w", net.branch_v.weight), ("branch_v_b", net.branch_v.bias)]
    else:
        out += [("warp_w", net.gaussian_warp.weight), ("warp_b", net.gaussian_warp.bias)]
    if hasattr(net, "gaussian_rotation"):   # (DeformNetworkODE has no rotation / scaling heads)
        out += [("rot_w", net.gaussian_rotation.weight), ("rot_b", net.gaussian_rotation.bias),
                ("scale_w", net.gaussian_scaling.weight), ("scale_b", net.gaussian_scaling.bias)]
    return out


def _fill_ptr_struct(struct, named):
    for name, t in named:
        ptr = 0 if t is None else t.data_ptr()
        if name.startswith("linear_w"):
            struct.linear_w[int(name[8:])] = ptr
        elif name.startswith("linear_b"):
            struct.linear_b[int(name[8:])] = ptr
        else:
            setattr(struct, name, ptr)
    return struct


class _FusedDeform(torch.autograd.Function):
    """One fused launch.  mode "times": t holds T time values, each applied to all N points (outputs [T,N,.]);
    mode "rows": t holds one value per row (any non-zero stride; outputs [N,.])."""

    @staticmethod
    def forward(ctx, net, need_grad, mode, xyz, t, *params):
        lib = _lib.load()
        dev = xyz.device
        if dev.type != "cuda":
            raise RuntimeError("DeformNetwork: tensors must live on a CUDA device (there is no CPU path)")
        xyz_c = xyz.detach().contiguous().float()
        rows = int(xyz_c.shape[0])
        t2 = t.detach()
        if t2.dtype != torch.float32:
            t2 = t2.float()
        cfg = net._cfg()
        wsb = C.c_size_t()
        with torch.cuda.device(dev):
            packed = net._packed(dev)
            stream = torch.cuda.current_stream().cuda_stream
            pstruct = _fill_ptr_struct(_lib.MlpParams(), _param_list(net))
            if mode == "times":
                t2 = t2.reshape(-1).contiguous()
                T = int(t2.numel())
                lead = (T, rows)
                _lib.check(lib.d3g_mlp_workspace_bytes_times(C.byref(cfg), rows, T, int(need_grad), C.byref(wsb)))
            else:
                T = 0
                lead = (rows,)
                if t2.dim() != 1 or t2.shape[0] != rows:
                    raise RuntimeError("DeformNetwork: per-row time tensor must have one value per point")
                _lib.check(lib.d3g_mlp_workspace_bytes(C.byref(cfg), rows, int(need_grad), C.byref(wsb)))
            ws = torch.empty(wsb.value, dtype=torch.uint8, device=dev)
            d_xyz = torch.empty(lead + ((4, 4) if net.is_6dof else (3,)), dtype=torch.float32, device=dev)
            d_rot = torch.empty(lead + (4,), dtype=torch.float32, device=dev) if net._heads & HEAD_ROT else None
            d_scale = torch.empty(lead + (3,), dtype=torch.float32, device=dev) if net._heads & HEAD_SCALE else None
            if mode == "times":
                _lib.check(lib.d3g_mlp_forward_times(C.byref(cfg), _lib.ptr(packed), C.byref(pstruct),
                                                     C.c_int64(rows), T, _lib.ptr(xyz_c), _lib.ptr(t2), _lib.ptr(ws),
                                                     int(need_grad), _lib.ptr(d_xyz), _lib.ptr(d_rot),
                                                     _lib.ptr(d_scale), C.c_void_p(stream)))
            else:
                _lib.check(lib.d3g_mlp_forward(C.byref(cfg), _lib.ptr(packed), C.byref(pstruct), C.c_int64(rows),
                                               _lib.ptr(xyz_c), _lib.ptr(t2), C.c_int64(int(t2.stride(0))),
                                               _lib.ptr(ws), int(need_grad), _lib.ptr(d_xyz), _lib.ptr(d_rot),
                                               _lib.ptr(d_scale), C.c_void_p(stream)))
            net._watch_status(ws)
        ctx.net, ctx.ws, ctx.rows, ctx.mode, ctx.T = net, ws, rows, mode, T
        ctx.xyz_requires_grad = xyz.requires_grad
        ctx.t_requires_grad = t.requires_grad
        ctx.save_for_backward(xyz_c, t2)
        ctx.set_materialize_grads(False)
        outs = [d_xyz]
        if d_rot is not None:
            outs.append(d_rot)
        if d_scale is not None:
            outs.append(d_scale)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gouts):
        lib = _lib.load()
        net = ctx.net
        xyz_c, t2 = ctx.saved_tensors
        dev = xyz_c.device
        rows = ctx.rows
        if ctx.ws is None:
            raise RuntimeError("DeformNetwork: backward through the same forward twice (the saved activations are "
                               "released after the first backward, as with retain_graph=False)")
        it = iter(gouts)
        g_xyz = next(it)
        g_rot = next(it) if net._heads & HEAD_ROT else None
        g_scale = next(it) if net._heads & HEAD_SCALE else None
        f = lambda g: None if g is None else g.contiguous().float()
        g_xyz, g_rot, g_scale = f(g_xyz), f(g_rot), f(g_scale)
        named = _param_list(net)
        # The kernels only ever accumulate (red.global.add / +=) into the gradient buffers.  With
        # net.fused_grad_accumulation set and every parameter already holding a contiguous fp32 .grad, they
        # add straight into those tensors and autograd receives None (no zero-fill, no AccumulateGrad add
        # per parameter).  Otherwise: one zero-filled flat buffer, returned as per-parameter views.
        direct = net.fused_grad_accumulation and all(
            p.grad is not None and p.grad.dtype == torch.float32 and p.grad.is_contiguous() and p.grad.device == dev
            for _, p in named if p.requires_grad)
        with torch.cuda.device(dev):
            if direct:
                scratch = None
                grads = []
                for _, p in named:
                    if p.requires_grad:
                        grads.append(p.grad)
                    else:  # frozen parameter: its gradient goes to a scratch buffer nobody reads
                        if scratch is None:
                            scratch = torch.empty(max(q.numel() for _, q in named), dtype=torch.float32, device=dev)
                        grads.append(scratch[:p.numel()].view_as(p))
            else:
                sizes = [(p.numel() + 3) // 4 * 4 for _, p in named]  # keep every view 16-byte aligned
                flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
                grads, off = [], 0
                for (_, p), n in zip(named, sizes):
                    grads.append(flat[off:off + p.numel()].view_as(p))
                    off += n
            gstruct = _fill_ptr_struct(_lib.MlpGrads(), [(n, g) for (n, _), g in zip(named, grads)])
            pstruct = _fill_ptr_struct(_lib.MlpParams(), named)
            cfg = net._cfg()
            packed = net._packed(dev, reuse_in_capture=True)
            dxyz_in = torch.empty(rows, 3, dtype=torch.float32, device=dev) if ctx.xyz_requires_grad else None
            dt_in = None
            if ctx.t_requires_grad:
                dt_in = torch.empty(ctx.T if ctx.mode == "times" else rows, dtype=torch.float32, device=dev)
            stream = torch.cuda.current_stream().cuda_stream
            if ctx.mode == "times":
                _lib.check(lib.d3g_mlp_backward_times(C.byref(cfg), _lib.ptr(packed), C.byref(pstruct),
                                                      C.c_int64(rows), ctx.T, _lib.ptr(xyz_c), _lib.ptr(t2),
                                                      _lib.ptr(ctx.ws), _lib.ptr(g_xyz), _lib.ptr(g_rot),
                                                      _lib.ptr(g_scale), C.byref(gstruct), _lib.ptr(dxyz_in),
                                                      _lib.ptr(dt_in), C.c_void_p(stream)))
            else:
                _lib.check(lib.d3g_mlp_backward(C.byref(cfg), _lib.ptr(packed), C.byref(pstruct), C.c_int64(rows),
                                                _lib.ptr(xyz_c), _lib.ptr(t2), C.c_int64(int(t2.stride(0))),
                                                _lib.ptr(ctx.ws), _lib.ptr(g_xyz), _lib.ptr(g_rot),
                                                _lib.ptr(g_scale), C.byref(gstruct), _lib.ptr(dxyz_in),
                                                _lib.ptr(dt_in), C.c_void_p(stream)))
            net._watch_status(ctx.ws)
        _FusedDeform.last_ws = ctx.ws if net.debug else None  # parity inspection hook (tools/dbg_mlp_bwd.py)
        ctx.ws = None
        # heads the network never evaluates get no gradient at all (like autograd in the reference), so
        # Adam's weight decay does not touch them
        unused = set()
        if not net._heads & HEAD_ROT:
            unused |= {"rot_w", "rot_b"}
        if not net._heads & HEAD_SCALE:
            unused |= {"scale_w", "scale_b"}
        if direct:
            out = tuple(None for _ in named)
        else:
            out = tuple(None if n in unused else g for (n, _), g in zip(named, grads))
        return (None, None, None, dxyz_in, dt_in) + out


class _DeformBase(nn.Module):
    """Shared constructor: identical module tree to utils/time_utils.py:57-100 / :130-174."""

    _rot_scale_heads = True   # (DeformNetworkODE: warp head only)

    def __init__(self, D=8, W=256, input_ch=3, output_ch=59, multires=10, is_blender=False, is_6dof=False):
        super().__init__()
        self.D, self.W = D, W
        self.output_ch = output_ch
        self.multires = multires
        self.t_multires = 6 if is_blender else 10
        self.skips = [D // 2]
        time_input_ch = 1 + 2 * self.t_multires
        xyz_input_ch = 3 + 6 * multires
        self.input_ch = xyz_input_ch + time_input_ch
        if is_blender:
            self.time_out = 30
            self.timenet = nn.Sequential(nn.Linear(time_input_ch, 256), nn.ReLU(inplace=True),
                                         nn.Linear(256, self.time_out))
            first = xyz_input_ch + self.time_out
        else:
            first = self.input_ch
        self.linear = nn.ModuleList(
            [nn.Linear(first, W)] + [nn.Linear(W, W) if i not in self.skips else nn.Linear(W + first, W)
                                     for i in range(D - 1)])
        self.is_blender = is_blender
        self.is_6dof = is_6dof
        if is_6dof:
            self.branch_w = nn.Linear(W, 3)
            self.branch_v = nn.Linear(W, 3)
        else:
            self.gaussian_warp = nn.Linear(W, 3)
        if self._rot_scale_heads:
            self.gaussian_rotation = nn.Linear(W, 4)
            self.gaussian_scaling = nn.Linear(W, 3)
        self.debug = bool(int(os.environ.get("D3G_DEBUG", "0")))
        # opt-in: backward adds into existing .grad tensors in place (see _FusedDeform.backward); only valid
        # with loss.backward()-style accumulation, not with torch.autograd.grad()
        self.fused_grad_accumulation = False
        self._pack_cache = None
        self._status_pending, self._status_host, self._status_next = [], None, 0
        self._graph_status, self._graph_pool = [], None   # pinned status slots of launches captured into CUDA graphs

    def _apply(self, fn, *a, **kw):
        self.__dict__.pop("_named_cache", None)
        self._pack_cache = None
        return super()._apply(fn, *a, **kw)

    def __setattr__(self, name, value):
        if isinstance(value, (nn.Parameter, nn.Module)):
            self.__dict__.pop("_named_cache", None)
        super().__setattr__(name, value)

    # -- C-ABI plumbing --------------------------------------------------------------------------
    def _cfg(self):
        c = _lib.MlpConfig()
        c.is_blender, c.is_6dof, c.heads = int(self.is_blender), int(self.is_6dof), int(self._heads)
        c.D, c.W, c.multires = int(self.D), int(self.W), int(self.multires)
        return c

    def _packed(self, dev, reuse_in_capture=False):
        """Packed weight images, rebuilt only when a parameter changed (tensor version counters)."""
        lib = _lib.load()
        named = _param_list(self)
        key = tuple((p.data_ptr(), p._version) for _, p in named) + (str(dev),)
        capturing = torch.cuda.is_current_stream_capturing()
        if capturing and reuse_in_capture and self._pack_cache is not None:
            return self._pack_cache[1]         # backward of a forward captured (and packed) in the same graph
        if not capturing and self._pack_cache is not None and self._pack_cache[0] == key:
            return self._pack_cache[1]
        # (inside a CUDA-graph capture the pack kernel always becomes part of the graph: replays must see the
        # parameter values of THEIR time, not those of capture time)
        cfg = self._cfg()
        nb = C.c_size_t()
        _lib.check(lib.d3g_mlp_packed_bytes(C.byref(cfg), C.byref(nb)))
        buf = self._pack_cache[1] if self._pack_cache is not None and self._pack_cache[1].device == dev \
            else torch.empty(nb.value, dtype=torch.uint8, device=dev)
        for _, p in named:
            if p.device != dev or p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("DeformNetwork parameters must be contiguous fp32 tensors on %s" % dev)
        pstruct = _fill_ptr_struct(_lib.MlpParams(), named)
        _lib.check(lib.d3g_mlp_pack_weights(C.byref(cfg), C.byref(pstruct), _lib.ptr(buf),
                                            C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        self._pack_cache = (key, buf)
        return buf

    def _watch_status(self, ws):
        """The kernels' bounded mbarrier waits set a status word instead of hanging.  It is copied to pinned host
        memory on the stream right behind the launch and examined once that copy has completed — at the next
        call, or in check_status() — so a pipeline abort never goes unnoticed and the hot path never syncs.
        With debug set (D3G_DEBUG=1) the check is immediate."""
        if torch.cuda.is_current_stream_capturing():
            # the copy becomes a node of the graph; every replay refreshes the slot, check_status() reads it.
            # (pinned memory cannot be allocated during a capture: the slots come from a pool made by an eager call)
            if self._graph_pool is None or len(self._graph_status) >= self._graph_pool.numel():
                raise RuntimeError("DeformNetwork: run the step eagerly once before capturing it into a CUDA graph")
            slot = self._graph_pool[len(self._graph_status):len(self._graph_status) + 1]
            slot.copy_(ws[:4].view(torch.int32), non_blocking=True)
            self._graph_status.append(slot)
            return
        self._poll_status(wait=len(self._status_pending) >= 32)
        if self._status_host is None:
            self._status_host = torch.zeros(64, dtype=torch.int32).pin_memory()
            self._graph_pool = torch.zeros(256, dtype=torch.int32).pin_memory()
        i = self._status_next
        self._status_next = (i + 1) % 64
        slot = self._status_host[i:i + 1]
        slot.copy_(ws[:4].view(torch.int32), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._status_pending.append((ev, slot))
        if self.debug:
            self.check_status()

    def _poll_status(self, wait=False):
        keep = []
        for ev, slot in self._status_pending:
            if wait:
                ev.synchronize()
            if wait or ev.query():
                if int(slot[0]) != 0:
                    self._status_pending = []
                    raise _lib.D3GError("fused MLP kernel reported an internal pipeline timeout (status %d); its "
                                        "outputs are invalid" % int(slot[0]))
            else:
                keep.append((ev, slot))
        self._status_pending = keep

    def check_status(self):
        """Wait for every launch issued so far and raise if one of them aborted."""
        self._poll_status(wait=True)
        if self._graph_status:
            torch.cuda.current_stream().synchronize()
            for slot in self._graph_status:
                if int(slot[0]) != 0:
                    raise _lib.D3GError("fused MLP kernel (CUDA-graph replay) reported an internal pipeline timeout "
                                        "(status %d); its outputs are invalid" % int(slot[0]))

    def read_activation_masks(self, ws, rows):
        """Parity hook: [8, rows, 256] bool activation pattern saved by the last training-mode forward."""
        lib = _lib.load()
        cfg = self._cfg()
        out = np.zeros((8, rows, 256), dtype=bool)
        buf = np.zeros(rows * 8, dtype=np.uint32)
        for l in range(8):
            _lib.check(lib.d3g_mlp_read_masks(C.byref(cfg), _lib.ptr(ws), C.c_int64(rows), l,
                                              buf.ctypes.data_as(C.c_void_p),
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            bits = (buf.reshape(rows, 8, 1) >> np.arange(32, dtype=np.uint32).reshape(1, 1, 32)) & 1
            out[l] = bits.reshape(rows, 256).astype(bool)
        return out

    def _run(self, x, t):
        """forward(x, t) of the reference: t is [N,1] (utils/time_utils.py:102-103).  One value shared by every
        row (the `fid.expand(N,1)` of train.py:207, or a single element) goes down the time-block path with
        T = 1; anything else is a per-row launch, whatever its stride."""
        params = [p for _, p in _param_list(self)]
        # (grad mode is off inside autograd.Function.forward, so decide here whether to save activations)
        need_grad = torch.is_grad_enabled() and (any(p.requires_grad for p in params) or x.requires_grad
                                                 or t.requires_grad)
        N = x.shape[0]
        if t.dim() == 2:
            if t.shape[1] != 1:
                raise RuntimeError("DeformNetwork.forward expects t of shape [N,1] (got %s)" % (tuple(t.shape),))
            shared = t.shape[0] == 1 or (t.shape[0] == N and t.stride(0) == 0)
            tv = t[:1].reshape(1) if shared else t[:, 0]
        elif t.dim() <= 1:
            shared = t.numel() == 1 or (t.numel() == N and t.stride(0) == 0)
            tv = t.reshape(-1)[:1] if shared else t
        else:
            raise RuntimeError("DeformNetwork.forward expects t of shape [N,1]")
        if not shared and tv.shape[0] != N:
            raise RuntimeError("DeformNetwork.forward: t has %d rows, x has %d" % (tv.shape[0], N))
        if shared:
            outs = _FusedDeform.apply(self, need_grad, "times", x, tv, *params)
            return tuple(o[0] for o in outs)
        return _FusedDeform.apply(self, need_grad, "rows", x, tv, *params)

    def forward_times(self, x, t_values):
        """T time values applied to the same N points in ONE launch: the loop of scene/deform_model.py:266-272
        without its T launches.  Returns tensors shaped [T,N,.]."""
        params = [p for _, p in _param_list(self)]
        need_grad = torch.is_grad_enabled() and (any(p.requires_grad for p in params) or x.requires_grad
                                                 or t_values.requires_grad)
        return _FusedDeform.apply(self, need_grad, "times", x, t_values.reshape(-1), *params)


class DeformNetworkBaseline(_DeformBase):
    """utils/time_utils.py:56-127: returns (d_xyz, rotation, scaling)."""
    _heads = HEAD_WARP | HEAD_ROT | HEAD_SCALE

    def forward(self, x, t):
        d_xyz, rotation, scaling = self._run(x, t)
        return d_xyz, rotation, scaling


class DeformNetwork(_DeformBase):
    """utils/time_utils.py:129-201: same trunk, returns (d_xyz, 0, 0)."""
    _heads = HEAD_WARP

    def forward(self, x, t):
        (d_xyz,) = self._run(x, t)
        return d_xyz, 0, 0


class DeformNetworkODE(_DeformBase):
    """utils/time_utils.py:331-438: the right-hand side dx/dt = f(t, x) of the `ode_batch*` configurations
    (scene/deform_model.py:183-207 hands it to odeint_adjoint).  With use_linear == 0 and use_emb (every shipped
    configuration, arguments/ode_*.py) it is the DeformNetwork trunk + gaussian_warp head, evaluated by the fused
    kernels; gradients flow to x, t and the parameters (the adjoint needs all three).  The use_linear 1..4 variants are
    single nn.Linear maps of the embedded input and stay torch ops, as in the reference."""
    _heads = HEAD_WARP
    _rot_scale_heads = False

    def __init__(self, D=8, W=256, input_ch=3, output_ch=59, multires=10, is_blender=False, is_6dof=False,
                 use_linear=0, use_emb=True, output_scale=1, skips=[4]):
        if use_linear == 0:
            if not use_emb or list(skips if skips is not None else []) != [4] or D != 8 or W != 256 or is_6dof:
                raise NotImplementedError(
                    "DeformNetworkODE: the fused kernels implement use_linear=0 with use_emb=True, skips=[4], D=8, "
                    "W=256, is_6dof=False (every configuration under arguments/); got use_emb=%s skips=%s D=%d W=%d "
                    "is_6dof=%s" % (use_emb, skips, D, W, is_6dof))
            super().__init__(D=D, W=W, input_ch=input_ch, output_ch=output_ch, multires=multires,
                             is_blender=is_blender, is_6dof=False)
            self.use_linear, self.use_emb, self.output_scale = use_linear, use_emb, output_scale
            return
        nn.Module.__init__(self)
        self.use_linear, self.use_emb, self.output_scale = use_linear, use_emb, output_scale
        self.D, self.W, self.input_ch, self.output_ch = D, W, input_ch, output_ch
        self.t_multires = 6 if is_blender else 10
        self.multires = multires
        self.is_blender, self.is_6dof = is_blender, is_6dof
        time_ch = 1 + 2 * self.t_multires if use_emb else 1
        xyz_ch = 3 + 6 * multires if use_emb else 3
        self.time_input_ch, self.xyz_input_ch, self.total_input_ch = time_ch, xyz_ch, time_ch + xyz_ch
        if use_linear == 1:
            self.linear_layer = nn.Linear(self.total_input_ch, 3)
        elif use_linear == 2:
            self.A_t_net = nn.Linear(time_ch, xyz_ch * xyz_ch)
            self.b_t_net = nn.Linear(time_ch, xyz_ch)
        elif use_linear == 3:
            self.xyz_linear = nn.Linear(xyz_ch, 3)
        elif use_linear == 4:
            self.z_linear = nn.Linear(1, 1)
        else:
            raise ValueError("use_linear must be 0..4")

    @staticmethod
    def _embed(x, L):
        out = [x]
        for f in range(L):
            out += [torch.sin(x * 2.0 ** f), torch.cos(x * 2.0 ** f)]
        return torch.cat(out, -1)

    def forward(self, t, x):
        if self.use_linear == 0:
            (d_xyz,) = self._run(x, t.reshape(1, 1) if t.numel() == 1 else t)
            return d_xyz * self.output_scale if self.output_scale != 1 else d_xyz
        if self.use_emb:
            t_emb = self._embed(t.expand(x.shape[0], 1), self.t_multires)
            x_emb = self._embed(x, self.multires)
        else:
            x_emb, t_emb = x, t.unsqueeze(1)
        if self.use_linear == 1:
            return self.linear_layer(torch.cat([x_emb, t_emb], dim=-1)) * self.output_scale
        if self.use_linear == 2:
            A_t = self.A_t_net(t_emb).view(-1, self.xyz_input_ch, self.xyz_input_ch)
            return (torch.bmm(A_t, x_emb.unsqueeze(-1)).squeeze(-1) + self.b_t_net(t_emb)) * self.output_scale
        if self.use_linear == 3:
            return self.xyz_linear(x_emb) * self.output_scale
        z = self.z_linear(x_emb[:, 2:3])
        return torch.cat([torch.zeros_like(x_emb[:, :2]), z], dim=1) * self.output_scale


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """utils/general_utils.py:78-109 (log-linear interpolation with optional warm-up)."""

    def helper(step):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        if lr_delay_steps > 0:
            delay_rate = lr_delay_mult + (1 - lr_delay_mult) * np.sin(0.5 * np.pi * np.clip(step / lr_delay_steps, 0, 1))
        else:
            delay_rate = 1.0
        tt = np.clip(step / max_steps, 0, 1)
        log_lerp = np.exp(np.log(lr_init) * (1 - tt) + np.log(lr_final) * tt)
        return delay_rate * log_lerp

    return helper


def _search_max_iteration(folder):
    return max(int(f.split("_")[-1]) for f in os.listdir(folder))


class _DeformModelBase:
    def save_weights(self, model_path, iteration):
        out = os.path.join(model_path, "deform/iteration_{}".format(iteration))
        os.makedirs(out, exist_ok=True)
        torch.save(self.deform.state_dict(), os.path.join(out, "deform.pth"))

    def load_weights(self, model_path, iteration=-1):
        it = _search_max_iteration(os.path.join(model_path, "deform")) if iteration == -1 else iteration
        path = os.path.join(model_path, "deform/iteration_{}/deform.pth".format(it))
        self.deform.load_state_dict(torch.load(path))

    def update_learning_rate(self, iteration):
        for group in self.optimizer.param_groups:
            if group["name"] == "deform":
                lr = self.deform_scheduler_args(iteration)
                group["lr"] = lr
                return lr


class DeformModel(_DeformModelBase):
    """scene/deform_model.py:253-308."""

    def __init__(self, is_blender=False, is_6dof=False, D=8, W=256, input_ch=3, output_ch=59, multires=10,
                 scale_lr=False):
        self.deform = DeformNetwork(is_blender=is_blender, is_6dof=is_6dof, D=D, W=W, input_ch=input_ch,
                                    output_ch=output_ch, multires=multires).cuda()
        self.optimizer = None
        self.spatial_lr_scale = 5
        self.scale_lr = scale_lr

    def step(self, xyz, time_emb):
        T = time_emb.shape[1]
        if time_emb.shape[0] == 1 or time_emb.stride(0) == 0:
            # every column is one time value (an expanded fid vector): all T timestamps in one launch
            (d_xyz,) = self.deform.forward_times(xyz, time_emb[0])
            return d_xyz, [0] * T, [0] * T
        d_xyz_list, d_rotation_list, d_scaling_list = [], [], []
        for i in range(T):
            d_xyz, d_rotation, d_scaling = self.deform(xyz, time_emb[:, i].unsqueeze(1))
            d_xyz_list.append(d_xyz)
            d_rotation_list.append(d_rotation)
            d_scaling_list.append(d_scaling)
        return torch.stack(d_xyz_list, dim=0), d_rotation_list, d_scaling_list

    def train_setting(self, training_args):
        l = [{"params": list(self.deform.parameters()),
              "lr": training_args.position_lr_init * self.spatial_lr_scale, "name": "deform"}]
        self.optimizer = _FusedAdam(l, lr=0.0, eps=1e-15, weight_decay=training_args.weight_decay)
        if self.scale_lr:
            lr_init = training_args.position_lr_init * self.spatial_lr_scale * training_args.num_cams_per_iter
            lr_final = training_args.position_lr_final * training_args.num_cams_per_iter
        else:
            lr_init = training_args.position_lr_init * self.spatial_lr_scale
            lr_final = training_args.position_lr_final
        self.deform_scheduler_args = get_expon_lr_func(lr_init=lr_init, lr_final=lr_final,
                                                       lr_delay_mult=training_args.position_lr_delay_mult,
                                                       max_steps=training_args.position_lr_max_steps)


class DeformModelBaseline(_DeformModelBase):
    """scene/deform_model.py:317-356."""

    def __init__(self, is_blender=False, is_6dof=False, D=8, W=256, input_ch=3, output_ch=59, multires=10):
        self.deform = DeformNetworkBaseline(is_blender=is_blender, is_6dof=is_6dof, D=D, W=W, input_ch=input_ch,
                                            output_ch=output_ch, multires=multires).cuda()
        self.optimizer = None
        self.spatial_lr_scale = 5

    def step(self, xyz, time_emb):
        return self.deform(xyz, time_emb)

    def train_setting(self, training_args):
        l = [{"params": list(self.deform.parameters()),
              "lr": training_args.position_lr_init * self.spatial_lr_scale, "name": "deform"}]
        self.optimizer = _FusedAdam(l, lr=0.0, eps=1e-15)
        self.deform_scheduler_args = get_expon_lr_func(
            lr_init=training_args.position_lr_init * self.spatial_lr_scale,
            lr_final=training_args.position_lr_final, lr_delay_mult=training_args.position_lr_delay_mult,
            max_steps=training_args.deform_lr_max_steps)


class DeformModelODE(_DeformModelBase):
    """scene/deform_model.py:172-251: positions at the sampled times are the solution of dx/dt = deform(t, x) from
    the canonical positions (adaptive dopri5 + adjoint).  torchdiffeq.odeint_adjoint is used when it is installed;
    otherwise this package's own solver of the same family (ode.py)."""

    def __init__(self, is_blender=False, is_6dof=False, D=8, W=256, input_ch=3, output_ch=59, multires=10,
                 scale_lr=False, use_linear=0, use_emb=True, rtol=0.001, atol=0.0001, output_scale=1, skips=[4]):
        self.deform = DeformNetworkODE(is_blender=is_blender, is_6dof=is_6dof, D=D, W=W, input_ch=input_ch,
                                       output_ch=output_ch, multires=multires, use_linear=use_linear, use_emb=use_emb,
                                       output_scale=output_scale, skips=skips).cuda()
        self.optimizer = None
        self.spatial_lr_scale = 5
        self.scale_lr = scale_lr
        self.rtol, self.atol = rtol, atol

    def step(self, xyz, time_emb):
        is_val = len(time_emb) == 1
        if is_val and float(time_emb[0]) == 0.0:
            return xyz, torch.zeros([xyz.shape[0], 4]).to(xyz.device), torch.zeros([xyz.shape[0], 3]).to(xyz.device)
        t_interval = torch.Tensor(time_emb).to(xyz.device)
        try:
            from torchdiffeq import odeint_adjoint
        except ImportError:
            from .ode import odeint_adjoint
        ode_value = odeint_adjoint(self.deform, xyz, t_interval, rtol=self.rtol, atol=self.atol)
        xyz_new = torch.squeeze(ode_value)
        rotation = torch.zeros([xyz_new.shape[0], xyz_new.shape[1], 4]).to(xyz.device)
        scale = torch.zeros([xyz_new.shape[0], xyz_new.shape[1], 3]).to(xyz.device)
        return xyz_new, rotation, scale

    def train_setting(self, training_args):
        l = [{"params": list(self.deform.parameters()),
              "lr": training_args.position_lr_init * self.spatial_lr_scale, "name": "deform"}]
        self.optimizer = _FusedAdam(l, lr=0.0, eps=1e-15, weight_decay=training_args.weight_decay)
        if self.scale_lr:
            lr_init = training_args.position_lr_init * self.spatial_lr_scale * training_args.num_cams_per_iter
            lr_final = training_args.position_lr_final * self.spatial_lr_scale * training_args.num_cams_per_iter
        else:
            lr_init = training_args.position_lr_init * self.spatial_lr_scale
            lr_final = training_args.position_lr_final
        self.deform_scheduler_args = get_expon_lr_func(lr_init=lr_init, lr_final=lr_final,
                                                       lr_delay_mult=training_args.position_lr_delay_mult,
                                                       max_steps=training_args.position_lr_max_steps)
