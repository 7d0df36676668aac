"""Fused Adam for the two optimizers of the training step: `GaussianModel.optimizer` (scene/gaussian_model.py:127-136,
six groups, lr per group, eps 1e-15) and `DeformModel.optimizer` (scene/deform_model.py:275-295,326-337, optional
weight_decay).  A subclass of torch.optim.Adam, so construction, `param_groups`, the lr schedulers that write
`group['lr']`, `state_dict()` / `load_state_dict()` and the reference's densification surgery on
`optimizer.state[p]['exp_avg' | 'exp_avg_sq']` (scene/gaussian_model.py:242-307) all work unchanged; only `step()` is
replaced: ONE kernel launch (csrc/adam.cu) instead of torch's per-op multi-tensor sequence.  CUDA fp32 only."""
import ctypes as C

import torch

from . import _lib


class Adam(torch.optim.Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        by_cfg = {}
        for group in self.param_groups:
            if group.get("amsgrad") or group.get("maximize"):
                raise NotImplementedError("fused Adam: amsgrad / maximize are not used by the reference")
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.device.type != "cuda" or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("fused Adam: parameters must be contiguous fp32 CUDA tensors (no CPU path)")
                if p.grad.is_sparse:
                    raise RuntimeError("fused Adam does not support sparse gradients")
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                st = self.state[p]
                if len(st) == 0:   # same layout as torch.optim.Adam's lazy state
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                t = _lib.AdamTensor(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                                    p.numel(), float(group["lr"]), float(group["weight_decay"]), int(st["step"].item()))
                by_cfg.setdefault((float(beta1), float(beta2), float(group["eps"]), p.device.index), []).append((t, g, p))
        for (beta1, beta2, eps, dev), items in by_cfg.items():
            arr = (_lib.AdamTensor * len(items))(*[t for t, _, _ in items])
            with torch.cuda.device(dev):
                _lib.check(lib.d3g_adam_step(len(items), arr, C.c_double(beta1), C.c_double(beta2), C.c_double(eps),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            # the kernel wrote through raw pointers: tell autograd (and every cache keyed on tensor versions,
            # e.g. DeformNetwork's packed weight images) that the parameters changed
            for _, _, p in items:
                torch.autograd.graph.increment_version(p)
        return loss
