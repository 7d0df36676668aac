#!/usr/bin/env python
"""bench.py — deform+raster fwd+bwd frames/sec (BASELINE.json metric) on 1..8 B200.

One "step" = one batch of (camera, timestamp) frames through the reference-facing API of this repo:
    d_xyz, d_rot, d_scale = DeformModelBaseline.step(xyz.detach(), fid.expand(N,1))   # fused tcgen05 MLP
    out = render(cam, gaussians, pipe, bg, d_xyz, d_rot, d_scale)                     # CUDA rasterizer
    loss = 0.8*l1(out["render"], gt) + 0.2*(1-ssim(out["render"], gt)); loss.backward()   # raster bwd + MLP bwd
(scene/deform_model.py:323-324, gaussian_renderer/__init__.py:32-133, train_baseline.py:109-128).

Configurations (BASELINE.json `configs`, --config):
  C2 (default, the headline): 100k Gaussians, SH degree 3, 800x800, ONE frame per GPU per step (weak scaling);
     with N > 1 GPUs every rank renders its own frame and the Gaussian + MLP gradients are summed with one NCCL
     all-reduce of a flat buffer.
  C1 / C4: 50k @400x400, 200k 6-DoF @800x800, same step.
  C3: 300k Gaussians, a 10-frame camera/time batch per step (the per-iteration loop of train.py:148-154,224-250 with its
     cumulative-L1 loss weights), frames k = rank, rank+G, ... on each of G GPUs, all of a rank's timestamps in ONE MLP
     launch, one all-reduce per step (strong scaling over the batch; bounded by ceil(10/G) frames on the busiest rank).
  C5: render-only, 1M Gaussians @1920x1080, 100-timestamp sweep with a fixed camera under no_grad
     (render_baseline.py:57-74, render.py:186-202), timestamps k = rank mod G, no collective.
`--impl reference` times the reference's CPU implementation of the same path on the host cores (see CpuArm).
"""
import argparse
import importlib
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "deformable-3d-gaussians_b200"

CONFIGS = {
    # frames: frames per step (None = one per GPU, weak scaling); train: fwd+bwd + gradient exchange
    "C1": dict(N=50000, H=400, W=400, is_6dof=False, kind="blob", frames=None, train=True),
    "C2": dict(N=100000, H=800, W=800, is_6dof=False, kind="uniform", frames=None, train=True),
    "C3": dict(N=300000, H=800, W=800, is_6dof=False, kind="uniform", frames=10, train=True),
    "C4": dict(N=200000, H=800, W=800, is_6dof=True, kind="uniform", frames=None, train=True),
    "C5": dict(N=1000000, H=1080, W=1920, is_6dof=False, kind="uniform", frames=100, train=False),
    # diagnostic only: a frame so small that the step time is the host-side launch overhead of the path
    "host": dict(N=2000, H=64, W=64, is_6dof=False, kind="uniform", frames=None, train=True),
}
METRIC = "deform+raster fwd+bwd frames/sec @100k Gaussians 800x800, 1/2/4/8 B200"
LAMBDA_DSSIM = 0.2   # arguments/__init__.py:119


def pkg(sub):
    return importlib.import_module(PKG + "." + sub)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--eager", action="store_true", help="launch every step eagerly instead of replaying a CUDA graph")
    a = ap.parse_args()
    single = CONFIGS[a.config]["frames"] is None
    if a.steps is None:
        a.steps = 200 if single else (10 if a.config == "C3" else 3)   # 20 warm-up + 200 timed
    if a.warmup is None:
        a.warmup = 20 if single else 3
    return a


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=float(p["hbm_gbs"]), bf16_tflops=float(p["bf16_tflops"]),
                    bf16_tflops_sustained=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
                     "hw_thermal_slowdown": 0x40, "hw_power_brake": 0x80}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.02)
        except Exception as e:  # NVML missing: report that rather than fail the bench
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def frame_schedule(cfg, world):
    """(theta_deg, fid) of every frame of one step, in time order."""
    T = cfg["frames"] if cfg["frames"] is not None else world
    if not cfg["train"]:   # C5: fixed camera, t = i/(T-1) (render.py:186-202)
        return [(20.0, k / (T - 1)) for k in range(T)]
    if T == 1:
        return [(20.0, 0.37)]
    return [(-180.0 + 360.0 * (k + 0.5) / T, k / (T - 1)) for k in range(T)]


def scaled_network(deform, cfg, device=None):
    """DeformNetworkBaseline(is_blender=True) with torch's seeded default init, heads scaled so the splats stay sane
    (SURVEY.md 8(d)); both bench arms use exactly these parameters."""
    import torch
    torch.manual_seed(0)
    model = deform.DeformModelBaseline(is_blender=True, is_6dof=cfg["is_6dof"])
    with torch.no_grad():
        net = model.deform
        for m in (net.gaussian_rotation, net.gaussian_scaling):
            m.weight.mul_(0.01); m.bias.mul_(0.01)
        if cfg["is_6dof"]:
            for m in (net.branch_w, net.branch_v):
                m.weight.mul_(0.01); m.bias.mul_(0.01)
        else:
            net.gaussian_warp.weight.mul_(0.1); net.gaussian_warp.bias.mul_(0.1)
    return model


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
class Workload:
    """Scene, network and this rank's frames of one step on `device`; host copies of the per-frame inputs pinned."""

    def __init__(self, cfg, rank, world, device):
        import torch
        synth, deform, frames = pkg("synth"), pkg("deform"), pkg("frames")
        self.torch = torch
        self.cfg, self.device, self.world = cfg, device, world
        N, H, W = cfg["N"], cfg["H"], cfg["W"]
        g = synth.SynthGaussians(N, seed=0, kind=cfg["kind"], sh_degree=3, device="cpu")
        for name in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
            setattr(g, name, getattr(g, name).to(device).requires_grad_(cfg["train"]))
        self.gaussians = g
        self.model = scaled_network(deform, cfg)
        sched = frame_schedule(cfg, world)
        self.frames_per_step = len(sched)
        self.mine = frames.shard_frames(len(sched), world, rank)
        self.frames_per_rank = [len(frames.shard_frames(len(sched), world, r)) for r in range(world)]
        weights = frames.frame_loss_weights(len(sched), LAMBDA_DSSIM) if cfg["frames"] else [(1.0 - LAMBDA_DSSIM, LAMBDA_DSSIM)]
        self.cams_host, self.cams, self.host, self.dev, self.weights = [], [], [], [], []
        for j, k in enumerate(self.mine):
            theta, fid = sched[k]
            cam = synth.make_camera(W, H, theta_deg=theta, fid=fid)
            self.cams_host.append(cam)
            h = dict(view=cam.world_view_transform.pin_memory(), proj=cam.full_proj_transform.pin_memory(),
                     campos=cam.camera_center.pin_memory(), fid=cam.fid.pin_memory())
            if cfg["train"]:
                h["gt"] = torch.rand(3, H, W, generator=torch.Generator().manual_seed(100 + k)).pin_memory()
            self.host.append(h)
            self.dev.append({kk: v.to(device) for kk, v in h.items()})
            self.cams.append(synth.camera_to(cam, device))
            self.weights.append(weights[k if cfg["frames"] else 0])
        self.h2d_bytes = sum(t.numel() * t.element_size() for h in self.host for t in h.values())
        self.fids_dev = torch.zeros(max(len(self.mine), 1), device=device)
        self.bg = torch.zeros(3, device=device)
        self.pipe = type("Pipe", (), dict(convert_SHs_python=False, compute_cov3D_python=False, debug=False))()
        if cfg["train"]:
            # flat gradient buffer: parameter .grad tensors are views into it, so the multi-GPU exchange is
            # one all-reduce with no packing kernel
            # world > 1: the Gaussian-parameter slice (92 % of the buffer) is all-reduced from inside backward, as soon as
            # the rasterizer's backward has accumulated it, and overlaps the MLP backward; the persistent MLP kernels
            # leave D3G_BENCH_SM_RESERVE SMs to the NCCL kernel (they otherwise own every SM and NCCL waits for them)
            self.early = world > 1 and os.environ.get("D3G_BENCH_EARLY", "1") == "1"
            self.grads = frames.FlatGrads(g.parameters() + list(self.model.deform.parameters()),
                                          early=len(g.parameters()) if self.early else 0)
            if self.early:
                pkg("_lib").load().d3g_mlp_sm_reserve(int(os.environ.get("D3G_BENCH_SM_RESERVE", "8")))
            # one backward() covers all of this rank's frames (their losses are summed first), so every parameter's
            # gradient is accumulated once per step: the early slice is complete after ONE accumulation per parameter
            self.grads.frames_per_reduce = 1
            self.flat = self.grads.flat
            # MLP backward adds straight into those .grad views (no per-parameter zero-fill + AccumulateGrad add)
            self.model.deform.fused_grad_accumulation = True
            # ... and so do the rasterizer / prepare backward kernels for the six Gaussian parameters: no `get_features`
            # concatenation, no six AccumulateGrad adds per frame.  Their hooks no longer fire, so the early all-reduce
            # is started by the renderer's per-frame callback instead.
            self.grad_sinks = os.environ.get("D3G_BENCH_GRAD_SINKS", "1") == "1"
            if self.grad_sinks:
                renderer = pkg("renderer")
                renderer.FUSED_GRAD_ACCUMULATION = True
                self.grads.frames_per_exchange = max(len(self.mine), 1)
                renderer.ON_GAUSSIAN_GRADS = self.grads.on_frame_grads if self.early else None
        else:
            self.grads = self.flat = None
            self.grad_sinks = False
        # sync-free frames: the binning workspace is sized from the largest pair count seen so far (+ headroom) instead
        # of reading `num_rendered` back in the middle of every frame; overflow is checked where the host syncs anyway
        ras = pkg("rasterizer")
        self.capacity = ras.GrowingPairCapacity(1.1 if len(sched) == world else 1.5)
        ras.PAIR_CAPACITY_POLICY = self.capacity
        self.image_loss = pkg("loss_utils")._ImageLoss
        self.loss_host = torch.zeros(8).pin_memory()
        self.steps_done = 0
        self.losses = []
        self.copy_stream = torch.cuda.Stream(device=device)
        self.gt_ready = torch.cuda.Event()
        self.last = None

    def _upload(self):
        torch = self.torch
        for cam, h, d in zip(self.cams, self.host, self.dev):
            cam.world_view_transform.copy_(h["view"], non_blocking=True)
            cam.full_proj_transform.copy_(h["proj"], non_blocking=True)
            cam.camera_center.copy_(h["campos"], non_blocking=True)
            cam.fid.copy_(h["fid"], non_blocking=True)
        if self.cfg["train"]:
            # the ground-truth images are only needed by the loss: upload them on a copy stream so the PCIe transfer
            # overlaps the deformation MLP and the rasterizer of the same step
            main = torch.cuda.current_stream()
            self.copy_stream.wait_stream(main)  # last step's loss has finished reading the buffers
            with torch.cuda.stream(self.copy_stream):
                for h, d in zip(self.host, self.dev):
                    d["gt"].copy_(h["gt"], non_blocking=True)
                self.gt_ready.record(self.copy_stream)

    def step(self, from_host):
        """One step = this rank's frames, fwd (+bwd).  from_host: copy the step's inputs from pinned host memory first
        and read the result back (the end-to-end leg); otherwise everything is already resident."""
        torch = self.torch
        render = pkg("renderer").render
        g, cfg = self.gaussians, self.cfg
        if from_host:
            self._upload()
        if not self.mine:
            return None
        N = g.get_xyz.shape[0]
        T = len(self.mine)
        if cfg["train"]:
            self.flat.zero_()
            if T == 1:   # the reference's single-frame call (train_baseline.py:109-115)
                time_input = self.cams[0].fid.unsqueeze(0).expand(N, -1)
                d_xyz, d_rot, d_scale = self.model.step(g.get_xyz.detach(), time_input)
                d_xyz, d_rot, d_scale = [d_xyz], [d_rot], [d_scale]
            else:        # every timestamp of this rank's share of the batch in ONE launch
                for j, cam in enumerate(self.cams):
                    self.fids_dev[j:j + 1].copy_(cam.fid, non_blocking=True)
                d_xyz, d_rot, d_scale = self.model.deform.forward_times(g.get_xyz.detach(), self.fids_dev[:T])
            if from_host:
                torch.cuda.current_stream().wait_event(self.gt_ready)
            loss = None
            for j, cam in enumerate(self.cams):
                out = render(cam, g, self.pipe, self.bg, d_xyz[j], d_rot[j], d_scale[j], cfg["is_6dof"])
                wl1, wss = self.weights[j]
                lj = self.image_loss.apply(out["render"], self.dev[j]["gt"], wl1, -wss, wss)  # train.py:245-250
                loss = lj if loss is None else loss + lj
            loss.backward()
            result = loss.detach()
            del loss, out, d_xyz, d_rot, d_scale   # (nothing may keep the autograd graph alive across steps)
        else:
            with torch.no_grad():
                result = torch.zeros((), device=self.device)
                CH = 4   # timestamps per MLP launch
                for c0 in range(0, T, CH):
                    idx = list(range(c0, min(T, c0 + CH)))
                    for j in idx:
                        self.fids_dev[j:j + 1].copy_(self.cams[j].fid, non_blocking=True)
                    d_xyz, d_rot, d_scale = self.model.deform.forward_times(g.get_xyz, self.fids_dev[c0:c0 + len(idx)])
                    for i, j in enumerate(idx):
                        out = render(self.cams[j], g, self.pipe, self.bg, d_xyz[i], d_rot[i], d_scale[i], False)
                        result = result + out["render"][:, ::64, ::64].sum()   # a checksum that depends on every frame
        if from_host:
            # the result travels to pinned host memory every step; the host reads it when it next waits for this
            # step (timed_region keeps at most two steps in flight), like a training loop that logs one step late
            slot = self.steps_done % self.loss_host.numel()
            self.loss_host[slot:slot + 1].copy_(result.reshape(1), non_blocking=True)
            self.steps_done += 1
            return slot
        return None

    def exchange(self):
        if self.world > 1 and self.cfg["train"]:
            self.grads.all_reduce()

    def enable_graphs(self):
        """Capture step(False) and step(True) into CUDA graphs (frames.GraphedStep); later step() calls replay them.
        The multi-GPU gradient exchange stays an eager NCCL call after the replay."""
        frames = pkg("frames")
        self._graphs = {}
        eager, eager_exchange = self.step, self.exchange
        in_graph = self.world > 1 and self.cfg["train"] and os.environ.get("D3G_BENCH_NCCL_IN_GRAPH", "1") == "1"

        def both(fh):
            slot = eager(fh)
            eager_exchange()
            return slot
        for from_host in (False, True):
            fn = (lambda fh=from_host: both(fh)) if in_graph else (lambda fh=from_host: eager(fh))
            self._graphs[from_host] = frames.GraphedStep(fn, warmup=2,
                                                         checks=[self.capacity.check, self.model.deform.check_status])
        self._eager_step = eager
        self.collective_in_graph = in_graph
        if in_graph:
            self.exchange = lambda: None     # the all-reduces are nodes of the step graph

        def replay(from_host):
            g = self._graphs[from_host]
            slot = g()
            if from_host:
                self.steps_done += 1
            return slot
        self.step = replay


def timed_region(work, steps, from_host, world, dist):
    """barrier + sync, K steps, sync + barrier; returns (max-over-ranks seconds, per-step ms list of this rank)."""
    torch = work.torch
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    # the host may run at most `depth` steps ahead of the device (launch latency stays hidden, the queues of
    # several ranks cannot drift arbitrarily far apart); 0 = unbounded
    depth = int(os.environ.get("D3G_BENCH_INFLIGHT", "2"))
    marks, stamps = [], [e0]

    def retire(block_all=False):
        while marks and (block_all or len(marks) >= max(depth, 1)):
            ev, slot = marks.pop(0)
            ev.synchronize()
            if slot is not None:   # end-to-end leg: consume the step's result on the host, verify its pair capacity
                work.losses.append(float(work.loss_host[slot]))
                work.capacity.check()

    for _ in range(steps):
        retire()
        slot = work.step(from_host)
        work.exchange()
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        stamps.append(ev)
        marks.append((ev, slot))
        if not depth:
            retire(block_all=from_host)
    retire(block_all=True)
    e1.record()
    torch.cuda.synchronize()
    work.capacity.check()
    sec = e0.elapsed_time(e1) / 1e3
    per_step = [stamps[i].elapsed_time(stamps[i + 1]) for i in range(len(stamps) - 1)]
    if world > 1:
        t = torch.tensor([sec], device=work.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
        dist.barrier()
    return sec, per_step


def kernel_table(work, steps):
    """Second pass with the library's per-kernel CUDA-event recorder on (events on the launching stream);
    returns {kernel: (avg ms per launch, launches per step)} and the mean pair count R of a frame."""
    import ctypes as C
    torch = work.torch
    lib = pkg("_lib").load()
    n = lib.d3g_prof_count()
    torch.cuda.synchronize()
    lib.d3g_prof_enable(1)
    step = getattr(work, "_eager_step", work.step)   # (the recorder brackets launches, so this pass runs eagerly)
    exchange = work.grads.all_reduce if (work.world > 1 and work.cfg["train"]) else (lambda: None)
    for _ in range(steps):
        step(False)
        exchange()
    ms = (C.c_double * n)()
    cnt = (C.c_int32 * n)()
    pkg("_lib").check(lib.d3g_prof_collect(ms, cnt, n))
    lib.d3g_prof_enable(0)
    table = {}
    for i in range(n):
        if cnt[i]:
            table[lib.d3g_prof_name(i).decode()] = (ms[i] / cnt[i], cnt[i] / steps)
    work.capacity.check()
    R = max(work.capacity.seen.values())
    return table, R


def ncu_traffic(config_name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of
    this config (profiles/traffic_<config>.json, written by tools/summarize_ncu.py traffic); {} if none."""
    path = os.path.join(ROOT, "profiles", "traffic_%s.json" % config_name)
    if not os.path.exists(path):
        return {}
    with open(path) as f:
        return json.load(f).get("kernels", {})


def roofline(table, R, cfg, pk, rows_per_mlp_launch, traffic=None):
    """Algorithmic bytes / flops per launch (DESIGN.md kernels table) over measured duration."""
    traffic = traffic or {}
    N, P = cfg["N"], cfg["H"] * cfg["W"]
    Tl = ((cfg["W"] + 15) // 16) * ((cfg["H"] + 15) // 16)
    R = R or 0
    # bytes each rasterizer stage must move once (inputs read once, outputs written once)
    algo_bytes = {
        "preprocess_fwd": 311 * N, "scan": 8 * N, "duplicate_with_keys": 20 * N + 12 * R, "radix_sort": 24 * R,
        "tile_sort": 24 * R, "tile_count": 12 * N + 4 * R, "tile_scatter": 28 * N + 12 * R, "tile_scan": 16 * Tl,
        "identify_tile_ranges": 8 * R + 8 * Tl, "render_fwd": 44 * R + 24 * P,
        "render_bwd": 88 * R + 32 * P, "preprocess_bwd": 600 * N,
    }
    # ALGORITHMIC flops per point (519,936 MAC forward incl. heads; dgrad chain 7*65536+16*256 MAC, wgrad = forward
    # MACs).  The kernels ISSUE more tensor work than that: three split-fp16 products in the forward, two in the dgrad
    # chain (kind::f16 at the bf16 rate, so the bf16 peak is the denominator); `issued` says how much.
    M = rows_per_mlp_launch
    mlp_flops = {"mlp_fwd": (2 * 519936 * M, 3.0), "mlp_dgrad": (2 * (7 * 65536 + 4096) * M, 2.0),
                 "mlp_wgrad": (2 * (7 * 65536 + 2 * 96 * 256 + 16 * 256) * M, 1.0)}
    rows = []
    for name, (ms, per_step) in table.items():
        if name in algo_bytes:
            ach = algo_bytes[name] / (ms * 1e-3) / 1e9
            rows.append(dict(kernel=name, bound="hbm", achieved=ach, peak=pk["hbm_gbs"], unit="GB/s",
                             frac=ach / pk["hbm_gbs"], ms=ms, launches_per_step=per_step))
        elif name in mlp_flops:
            fl, mult = mlp_flops[name]
            ach = fl / (ms * 1e-3) / 1e12
            rows.append(dict(kernel=name, bound="tensor", achieved=ach, peak=pk["bf16_tflops_sustained"], unit="TFLOP/s",
                             frac=ach / pk["bf16_tflops_sustained"], issued=ach * mult,
                             issued_frac=ach * mult / pk["bf16_tflops_sustained"], ms=ms, launches_per_step=per_step))
        else:
            rows.append(dict(kernel=name, bound=None, ms=ms, launches_per_step=per_step))
    for r in rows:
        r["traffic"] = traffic.get(r["kernel"])
    rows.sort(key=lambda r: -r["ms"] * r["launches_per_step"])
    return rows


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's CPU implementation of the path on the host cores
# ------------------------------------------------------------------------------------------------
class CpuArm:
    """One FULL frame of the configuration per step, nothing scaled:
      * deformation MLP: the reference's DeformNetworkBaseline.forward (utils/time_utils.py:102-127) as it runs —
        torch ops (nn.Linear, ReLU, cat, sin/cos), forward + autograd backward, torch.set_num_threads(all host cores).
        /root/reference does not exist on the GPU box, so the module is restated op by op (kind "port"); it carries
        the SAME parameters as the GPU arm (one state_dict).
      * rasterizer: the reference has no CPU rasterizer (its CUDA extension's source is absent from the archive); the
        C restatement under oracle/ (OpenMP over tiles) stands in, forward + backward, labelled "port".
      * loss: 0.8 L1 + 0.2 (1 - SSIM) of the reference's utils/loss_utils.py restated in numpy (oracle/loss_ref.py).
    """

    def __init__(self, cfg):
        import torch
        from oracle import raster_oracle as ro, loss_ref
        synth, deform = pkg("synth"), pkg("deform")
        self.torch, self.ro, self.loss_ref, self.cfgd = torch, ro, loss_ref, cfg
        N, H, W = cfg["N"], cfg["H"], cfg["W"]
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        g = synth.SynthGaussians(N, seed=0, kind=cfg["kind"], sh_degree=3)
        theta, fid = frame_schedule(cfg, 1)[0]
        cam = synth.make_camera(W, H, theta_deg=theta, fid=fid)
        self.g, self.fid = g, float(fid)
        self.rcfg = ro.make_cfg(H, W, math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), np.zeros(3), 1.0,
                                cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(),
                                cam.camera_center.numpy(), 3, 16, 2, np.float32)
        sd = {k: v.detach().cpu() for k, v in scaled_network(deform_cpu_shim(), cfg).deform.state_dict().items()}
        self.net = TorchDeform(cfg["is_6dof"])
        self.net.load_state_dict(sd)
        self.gt = torch.rand(3, H, W, generator=torch.Generator().manual_seed(100)).numpy()

    def frame(self):
        torch, cfg, g = self.torch, self.cfgd, self.g
        N = g.get_xyz.shape[0]
        xyz = g.get_xyz
        t = torch.full((1, 1), self.fid).expand(N, 1)
        d_xyz, d_rot, d_scale = self.net(xyz, t)
        with torch.no_grad():
            if cfg["is_6dof"]:
                hom = torch.cat([xyz, torch.ones(N, 1)], 1)
                m = torch.einsum("nij,nj->ni", d_xyz, hom)
                means = m[:, :3] / m[:, 3:]
            else:
                means = xyz + d_xyz
            scales = g.get_scaling + d_scale
            rots = g.get_rotation + d_rot
        ro = self.ro
        fwd = ro.forward(self.rcfg, means.numpy(), g.get_features.numpy(), None, g.get_opacity.numpy(), scales.numpy(),
                         rots.numpy(), None, np.float32)
        dimg = self.loss_ref.training_loss_grad(fwd["color"].astype(np.float32), self.gt, LAMBDA_DSSIM).astype(np.float32)
        gb = ro.backward(self.rcfg, fwd, dimg, None, np.float32)
        R = fwd["R"]
        gr = dict(means3D=gb["dL_dmeans3D"], rots=gb["dL_drots"], scales=gb["dL_dscales"])
        if cfg["is_6dof"]:
            gd = torch.zeros(N, 4, 4)   # (chain through from_homogenous omitted: same MLP backward cost)
            gd[:, :3, 3] = torch.from_numpy(gr["means3D"])
        else:
            gd = torch.from_numpy(gr["means3D"])
        torch.autograd.backward([d_xyz, d_rot, d_scale], [gd, torch.from_numpy(gr["rots"]), torch.from_numpy(gr["scales"])])
        for p in self.net.parameters():
            p.grad = None
        return R

    def time(self, steps, warmup, budget_s=None):
        """Returns (seconds per frame, steps actually timed, warm-up frames actually run)."""
        t0 = time.perf_counter()
        for _ in range(max(warmup, 1) if budget_s is None else 1):
            self.frame()
        w = max(warmup, 1) if budget_s is None else 1
        first = time.perf_counter() - t0
        if budget_s is not None:
            steps = max(1, min(steps, int((budget_s - first) / max(first / w, 1e-6))))
        t0 = time.perf_counter()
        for _ in range(steps):
            self.frame()
        return (time.perf_counter() - t0) / steps, steps, w

    def sample_text(self, steps):
        c = self.cfgd
        return ("%d full frame(s) of the configuration (%d Gaussians, %dx%d, nothing scaled): torch-CPU "
                "DeformNetworkBaseline fwd+bwd restated op by op from utils/time_utils.py:102-127 with the GPU arm's "
                "parameters + C/OpenMP rasterizer restatement fwd+bwd + numpy L1/SSIM loss" % (steps, c["N"], c["W"], c["H"]))


def deform_cpu_shim():
    """scaled_network() only needs a `DeformModelBaseline(is_blender, is_6dof)` factory whose `.deform` has the
    reference's parameter names; on the reference arm no CUDA library may be touched, so this builds the module tree of
    the product's DeformNetworkBaseline on the CPU without loading libd3g.so."""
    deform = pkg("deform")

    class _M:
        def __init__(self, is_blender=False, is_6dof=False):
            self.deform = deform.DeformNetworkBaseline(is_blender=is_blender, is_6dof=is_6dof)

    return type("shim", (), {"DeformModelBaseline": _M})


def TorchDeform(is_6dof):
    import torch
    import torch.nn as nn

    def embed(x, L):
        out = [x]
        for f in range(L):
            out += [torch.sin(x * 2.0 ** f), torch.cos(x * 2.0 ** f)]
        return torch.cat(out, -1)

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.timenet = nn.Sequential(nn.Linear(13, 256), nn.ReLU(inplace=True), nn.Linear(256, 30))
            self.linear = nn.ModuleList([nn.Linear(93, 256)] + [nn.Linear(256, 256) if i != 4 else nn.Linear(349, 256)
                                                                 for i in range(7)])
            if is_6dof:
                self.branch_w, self.branch_v = nn.Linear(256, 3), nn.Linear(256, 3)
            else:
                self.gaussian_warp = nn.Linear(256, 3)
            self.gaussian_rotation, self.gaussian_scaling = nn.Linear(256, 4), nn.Linear(256, 3)

        def forward(self, x, t):
            te = self.timenet(embed(t, 6))
            xe = embed(x, 10)
            h = torch.cat([xe, te], -1)
            for i, l in enumerate(self.linear):
                h = torch.relu(l(h))
                if i == 4:
                    h = torch.cat([xe, te, h], -1)
            if is_6dof:   # utils/time_utils.py:114-121 + utils/rigid_utils.py:60-83
                w, v = self.branch_w(h), self.branch_v(h)
                theta = torch.norm(w, dim=-1, keepdim=True)
                w, v = w / theta + 1e-5, v / theta + 1e-5
                W = torch.zeros(w.shape[0], 3, 3)
                W[:, 0, 1], W[:, 0, 2], W[:, 1, 0] = -w[:, 2], w[:, 1], w[:, 2]
                W[:, 1, 2], W[:, 2, 0], W[:, 2, 1] = -w[:, 0], -w[:, 1], w[:, 0]
                th = theta[..., None]
                I = torch.eye(3)[None]
                Rm = I + torch.sin(th) * W + (1 - torch.cos(th)) * (W @ W)
                p = (th * I + (1 - torch.cos(th)) * W + (th - torch.sin(th)) * (W @ W)) @ v[..., None]
                top = torch.cat([Rm, p], -1)
                d_xyz = torch.cat([top, torch.tensor([0., 0., 0., 1.]).expand(w.shape[0], 1, 4)], 1)
            else:
                d_xyz = self.gaussian_warp(h)
            return d_xyz, self.gaussian_rotation(h), self.gaussian_scaling(h)

    return Net()


def workload_text(name, cfg, world):
    if cfg["frames"] is None:
        return "%s: %d Gaussians SH3 %dx%d deform MLP + rasterize fwd+bwd, 1 frame per GPU per step" % (
            name, cfg["N"], cfg["W"], cfg["H"])
    if cfg["train"]:
        return ("%s: %d Gaussians SH3 %dx%d, %d-frame camera/time batch per step sharded over %d GPU(s), deform MLP "
                "(one launch per rank) + rasterize fwd+bwd, one gradient all-reduce" % (
                    name, cfg["N"], cfg["W"], cfg["H"], cfg["frames"], world))
    return "%s: render-only, %d Gaussians SH3 %dx%d, %d-timestamp sweep split over %d GPU(s), no_grad, no collective" % (
        name, cfg["N"], cfg["W"], cfg["H"], cfg["frames"], world)


def run_reference(args, cfg, rank):
    """--impl reference: rank 0 alone runs the CPU arm on the full configuration."""
    if rank != 0:
        return
    arm = CpuArm(cfg)
    sec, steps, warm = arm.time(args.steps, args.warmup, budget_s=float(os.environ.get("D3G_REF_BUDGET_S", "150")))
    fps = 1.0 / sec
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "steps_requested": args.steps, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_text(args.config, dict(cfg, frames=None), 1), "is_6dof": cfg["is_6dof"],
                       "note": "one frame per step on the host cores; steps are capped so that the run ends within "
                               "D3G_REF_BUDGET_S seconds and the steps actually timed are the ones reported"},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": arm.cores, "kind": "port",
                             "sample": arm.sample_text(steps)},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def pct(xs, q):
    s = sorted(xs)
    return s[min(len(s) - 1, max(0, int(round(q * (len(s) - 1)))))] if s else None


def main():
    args = parse_args()
    cfg = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, cfg, rank)
        return
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
        # a rank that blocks in a collective (a peer died, mismatched collectives) must not hold N GPUs until some outer
        # limit fires: the whole default run takes well under a minute
        wd = float(os.environ.get("D3G_BENCH_WATCHDOG_S", "240"))

        def _watchdog():
            print("bench.py: rank %d made no exit within %.0f s — aborting" % (rank, wd), file=sys.stderr, flush=True)
            os._exit(3)
        t = threading.Timer(wd, _watchdog)
        t.daemon = True
        t.start()
    work = Workload(cfg, rank, world, device)

    warm = max(args.warmup, 3)
    for _ in range(3):
        work.step(False)
        work.exchange()
    graphed = False
    if not args.eager and os.environ.get("D3G_BENCH_GRAPH", "1") == "1":
        try:
            work.enable_graphs()
            graphed = True
        except Exception as e:   # report and fall back to eager launches
            print("bench.py: CUDA-graph capture failed (%s: %s); running eagerly" % (type(e).__name__, e), file=sys.stderr)
            torch.cuda.synchronize()
    for _ in range(warm):
        work.step(False)
        work.exchange()
    sampler = ClockSampler(local_rank)
    sampler.start()
    sec, per_step = timed_region(work, args.steps, False, world, dist)
    work.step(True)  # warm the host-copy path
    torch.cuda.synchronize()
    work.losses.clear()
    sec_e2e, per_step_e2e = timed_region(work, args.steps, True, world, dist)
    sampler.stop_flag = True
    sampler.join(timeout=2.0)
    fps_step = work.frames_per_step
    value = fps_step * args.steps / sec
    e2e = fps_step * args.steps / sec_e2e

    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak" if cfg["frames"] is None else "strong", "vs_baseline": None,
            "dtype": "f16 split x3 (fp32 accumulate, fp32-class accuracy) MLP / f32 rasterizer", "data": "synthetic",
            "ms_per_step_percentiles": {"p10": pct(per_step, 0.1), "p50": pct(per_step, 0.5), "p90": pct(per_step, 0.9),
                                        "note": "this rank's per-step CUDA-event intervals inside the timed region"},
            "config": {"workload": workload_text(args.config, cfg, world), "is_6dof": cfg["is_6dof"],
                       "frames_per_step": fps_step, "frames_per_rank": work.frames_per_rank,
                       "busiest_rank_bound": ("ceil(%d/%d) = %d frames" % (fps_step, world, max(work.frames_per_rank))),
                       "net": "DeformNetworkBaseline(is_blender=True)",
                       "loss": ("0.8*L1 + 0.2*(1-SSIM) vs fixed random images (fused CUDA loss); multi-frame batches use "
                                "the reference's cumulative-L1 frame weights (train.py:245-250)") if cfg["train"] else None,
                       "pair_capacity": "GrowingPairCapacity: no num_rendered read-back inside a frame",
                       "grad_accumulation": ("backward kernels add into the flat .grad views (MLP and Gaussian parameters; "
                                             "SH coefficients read from _features_dc/_features_rest in place)"
                                             if getattr(work, "grad_sinks", False) else "autograd AccumulateGrad"),
                       "launch": ("one CUDA-graph replay per step (frames.GraphedStep: the whole fwd+bwd step incl. the "
                                  "H2D/D2H copies of the end-to-end leg is captured once)" if graphed else
                                  "eager: every kernel launched from Python each step"),
                       "l2": "per-step working set (saved MLP activation tiles, ~0.9 GB per 100k points) exceeds the 126 MB L2",
                       "parallelism": ("frame-parallel dp%d, NCCL all-reduce of the flat gradient buffer (Gaussian slice issued "
                                       "from inside backward, overlapping the MLP backward on reserved SMs: %s; collectives "
                                       "captured in the step graph: %s), <= 2 steps in flight" %
                                       (world, getattr(work, "early", False), getattr(work, "collective_in_graph", False)))
                       if cfg["train"] else ("timestamp-parallel x%d, no collective" % world)},
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": work.h2d_bytes, "d2h_bytes_per_step": 4,
                    "ms_per_step": sec_e2e / args.steps * 1e3, "results_read": len(work.losses),
                    "last_result": (work.losses[-1] if work.losses else None),
                    "p10_p50_p90_ms": [pct(per_step_e2e, 0.1), pct(per_step_e2e, 0.5), pct(per_step_e2e, 0.9)],
                    "pipelining": "inputs uploaded from pinned memory and the result copied to pinned memory every step; "
                                  "the host consumes step k's result while step k+1 runs (<= 2 steps in flight)"},
            "clocks": sampler.summary()}

    if not args.no_roofline:
        table, R = kernel_table(work, max(2, min(args.steps, 10)))
        rows_per_launch = cfg["N"] * (max(len(work.mine), 1) if cfg["train"] else min(4, max(len(work.mine), 1)))
        rows = roofline(table, R, cfg, peaks(), rows_per_launch, ncu_traffic(args.config))
        line["gpu_launches"] = int(round(sum(r["launches_per_step"] for r in rows) * args.steps))
        line["pairs_R"] = R
        top = next((r for r in rows if r.get("bound")), None)
        if top:
            line["roofline"] = {"kernel": top["kernel"], "bound": top["bound"], "achieved": top["achieved"],
                                "peak": top["peak"], "unit": top["unit"], "frac": top["frac"], "traffic": top["traffic"],
                                "note": ("render_* are FP32-issue bound, not HBM bound (ncu: sm__throughput ~80 %, DRAM "
                                         "< 1 %, profiles/); GB/s against the HBM peak is reported because the metric "
                                         "asks for it" if top["kernel"].startswith("render") else None),
                                "peak_source": peaks()["source"] + (" bf16 sustained (kind::f16 runs at the bf16 rate); "
                                                                    "tf32 dense measured 662 burst / 587 sustained "
                                                                    "(profiles/r2_numerics_split.json)"
                                                                    if top["bound"] == "tensor" else " copy bandwidth")}
        line["kernels"] = [{k: (round(v, 5) if isinstance(v, float) else v) for k, v in r.items()} for r in rows]
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        arm = CpuArm(cfg)
        csec, csteps, _ = arm.time(2, 1, budget_s=float(os.environ.get("D3G_CPU_BASELINE_BUDGET_S", "40")))
        line["cpu_baseline"] = {"value": 1.0 / csec, "unit": "frames/s", "cores": arm.cores, "kind": "port",
                                "sample": arm.sample_text(csteps)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        # tear down: graphs that captured NCCL kernels must go before the communicator; a watchdog ends the process if
        # the teardown itself blocks (the result line is already out)
        threading.Timer(20.0, lambda: os._exit(0)).start()
        torch.cuda.synchronize()
        dist.barrier()
        if hasattr(work, "_graphs"):
            work._graphs.clear()
            import gc
            gc.collect()
        torch.cuda.synchronize()
        dist.destroy_process_group()
        os._exit(0)


if __name__ == "__main__":
    main()
