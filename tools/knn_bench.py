"""distCUDA2 timing: ours (CUDA events, median of K launches) next to the reference compiled into oracle/_ref (host
clock around its synchronous call — it allocates, copies the bounding box to the host twice and syncs).
Usage: python tools/knn_bench.py [P ...]"""
import importlib
import os
import statistics
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
knn = importlib.import_module("deformable-3d-gaussians_b200.simple_knn")
from oracle import knn_oracle  # noqa: E402  (checker / baseline only)

for P in [int(a) for a in sys.argv[1:]] or [50000, 100000, 300000]:
    g = torch.Generator().manual_seed(0)
    pts = (torch.rand(P, 3, generator=g) * 2.6 - 1.3).cuda()
    line = "P=%d" % P
    for method in ("grid", "pairs"):
        if method == "pairs" and P > 300000:
            continue
        for _ in range(3):
            out = knn.distCUDA2(pts, method=method)
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); out = knn.distCUDA2(pts, method=method); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        line += " | ours[%s] %.3f ms" % (method, statistics.median(ts))
    if os.path.exists(knn_oracle.REF_SO):
        rs = []
        for _ in range(5):
            torch.cuda.synchronize(); t0 = time.perf_counter(); ref = knn_oracle.reference_mean_dist2(pts)
            rs.append((time.perf_counter() - t0) * 1e3)
        line += " | reference %.3f ms | bit-exact(last) %s" % (statistics.median(rs), bool(torch.equal(ref, out)))
    print(line, flush=True)
