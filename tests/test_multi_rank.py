"""Frame-sharded gradient exchange on 2 gloo ranks (CPU): summed gradients equal the serial loop's."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG, ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _frame_loss(params, k):
    w, b = params
    x = torch.full((4, 3), float(k + 1))
    return ((x @ w + b) ** 2).sum() * (k + 1)


def _worker(rank, world, port, T, out):
    import importlib
    import sys
    sys.path.insert(0, ROOT)
    frames = importlib.import_module(PKG + ".frames")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    params = [torch.randn(3, 2, requires_grad=True), torch.randn(2, requires_grad=True)]
    # the first parameter plays the role of the Gaussian slice: reduced asynchronously from inside the last
    # local frame's backward, the rest at the end
    fg = frames.FlatGrads(params, early=1)
    fg.frames_per_reduce = len(frames.shard_frames(T, world, rank))
    fg.zero()
    for k in frames.shard_frames(T, world, rank):
        _frame_loss(params, k).backward()
    assert params[0].grad.data_ptr() == fg.flat.data_ptr(), "grads must stay views of the flat buffer"
    fg.all_reduce()
    stats = [torch.tensor([1.0 + rank]), torch.tensor([2.0]), torch.tensor([float(rank * 5)])]
    frames.reduce_densification_stats(*stats)
    if rank == 0:
        torch.save({"flat": fg.flat.clone(), "grads": [p.grad.clone() for p in params], "stats": stats}, out)
    dist.destroy_process_group()


def test_two_rank_gradient_sum_matches_serial(tmp_path):
    T, world = 5, 2
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(world, _free_port(), T, out), nprocs=world, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    params = [torch.randn(3, 2, requires_grad=True), torch.randn(2, requires_grad=True)]
    sum(_frame_loss(params, k) for k in range(T)).backward()
    for g, p in zip(got["grads"], params):
        assert torch.allclose(g, p.grad, rtol=1e-6, atol=1e-6)
    # views are padded to 16-byte boundaries inside the flat buffer; the padding stays zero through the reduce
    assert got["flat"].numel() == 8 + 4 and got["flat"][6:8].abs().sum() == 0 and got["flat"][10:].abs().sum() == 0
    assert got["stats"][0].item() == 3.0 and got["stats"][1].item() == 4.0 and got["stats"][2].item() == 5.0


def _worker_summed(rank, world, port, T, out):
    """bench.py's shape: the losses of a rank's frames are summed and back-propagated ONCE per step (one gradient
    accumulation per parameter, whatever the rank's frame count), several steps in a row, uneven shards."""
    import importlib
    import sys
    sys.path.insert(0, ROOT)
    frames = importlib.import_module(PKG + ".frames")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    params = [torch.randn(3, 2, requires_grad=True), torch.randn(2, requires_grad=True)]
    fg = frames.FlatGrads(params, early=1)
    fg.frames_per_reduce = 1
    mine = frames.shard_frames(T, world, rank)
    results = []
    for step in range(3):
        fg.zero()
        if mine:
            sum(_frame_loss(params, k) for k in mine).backward()
        fg.all_reduce()
        results.append(fg.flat.clone())
    if rank == 0:
        torch.save(results, out)
    dist.destroy_process_group()


def test_uneven_shards_with_one_backward_per_step(tmp_path):
    T, world = 3, 2          # shards of 2 and 1 frames
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker_summed, args=(world, _free_port(), T, out), nprocs=world, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    params = [torch.randn(3, 2, requires_grad=True), torch.randn(2, requires_grad=True)]
    sum(_frame_loss(params, k) for k in range(T)).backward()
    for flat in got:      # every step reduces the same sums: no step may be skipped or doubled
        assert torch.allclose(flat[:6].view(3, 2), params[0].grad, rtol=1e-6, atol=1e-6)
        assert torch.allclose(flat[8:10], params[1].grad, rtol=1e-6, atol=1e-6)


def _worker_frame_callback(rank, world, port, T, out):
    """renderer.FUSED_GRAD_ACCUMULATION's shape: the early parameters' gradients are added by kernels (here: by hand,
    under no_grad — no AccumulateGrad, no hook) and the renderer reports each finished frame through
    FlatGrads.on_frame_grads(); the early slice is reduced after the rank's last frame, uneven shards included."""
    import importlib
    import sys
    sys.path.insert(0, ROOT)
    frames = importlib.import_module(PKG + ".frames")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    params = [torch.randn(3, 2, requires_grad=True), torch.randn(2, requires_grad=True)]
    fg = frames.FlatGrads(params, early=1)
    mine = frames.shard_frames(T, world, rank)
    fg.frames_per_exchange = len(mine)
    results = []
    for step in range(3):
        fg.zero()
        started = []
        for k in mine:
            gw, gb = torch.autograd.grad(_frame_loss(params, k), params)
            with torch.no_grad():
                params[0].grad += gw          # "kernel" accumulation of the early parameter
                params[1].grad += gb
            fg.on_frame_grads()
            started.append(fg._work is not None)
        assert started == [False] * (len(mine) - 1) + [True], started   # one early reduce, after the last frame
        fg.all_reduce()
        results.append(fg.flat.clone())
    if rank == 0:
        torch.save(results, out)
    dist.destroy_process_group()


def test_frame_callback_starts_the_early_reduce_once_per_step(tmp_path):
    T, world = 3, 2
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker_frame_callback, args=(world, _free_port(), T, out), nprocs=world, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    params = [torch.randn(3, 2, requires_grad=True), torch.randn(2, requires_grad=True)]
    sum(_frame_loss(params, k) for k in range(T)).backward()
    for flat in got:
        assert torch.allclose(flat[:6].view(3, 2), params[0].grad, rtol=1e-6, atol=1e-6)
        assert torch.allclose(flat[8:10], params[1].grad, rtol=1e-6, atol=1e-6)


def test_shard_and_weights():
    import importlib
    frames = importlib.import_module(PKG + ".frames")
    assert frames.shard_frames(10, 4, 1) == [1, 5, 9]
    assert sorted(sum((frames.shard_frames(10, 8, r) for r in range(8)), [])) == list(range(10))
    w = frames.frame_loss_weights(3, 0.2)
    assert abs(sum(a for a, _ in w) - (3 + 2 + 1) * 0.8 / 3) < 1e-12 and all(abs(b - 0.2 / 3) < 1e-12 for _, b in w)
