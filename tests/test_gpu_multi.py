"""GPU, >= 2 devices: frame-parallel data parallelism.  One process renders a 2-frame batch
sequentially and sums the gradients; two NCCL ranks render one frame each and all-reduce the flat gradient buffer.
The two flat buffers (Gaussian parameters + MLP parameters, 59 floats per Gaussian + 522,280) must agree within fp32
summation-order noise.  Run with `gpurun --gpus 2`; skipped on a single-GPU box."""
import os
import sys

import pytest
import torch

from conftest import ROOT, pkg

pytestmark = pytest.mark.gpu


def _batch_gradients(rank, world, frames_idx, out_path, port):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    bench = __import__("bench")
    torch.cuda.set_device(rank if world > 1 else 0)
    dev = torch.device("cuda", rank if world > 1 else 0)
    if world > 1:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cfg = dict(N=20000, H=256, W=256, is_6dof=False, kind="uniform", frames=2, train=True)
    work = bench.Workload(cfg, rank, world, dev)
    assert work.mine == frames_idx, (work.mine, frames_idx)
    work.step(False)
    work.exchange()
    torch.cuda.synchronize()
    work.capacity.check()
    if rank == 0:
        torch.save(work.flat.detach().cpu(), out_path)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _worker(rank, world, out_path, port):
    _batch_gradients(rank, world, [rank], out_path, port)


def test_two_ranks_reproduce_the_single_process_gradient_sum(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    one, two = str(tmp_path / "one.pt"), str(tmp_path / "two.pt")
    ctx = mp.get_context("spawn")
    p = ctx.Process(target=_batch_gradients, args=(0, 1, [0, 1], one, 0))
    p.start(); p.join()
    assert p.exitcode == 0
    mp.spawn(_worker, args=(2, two, 29533), nprocs=2, join=True)
    a, b = torch.load(one), torch.load(two)
    assert a.shape == b.shape and float(a.abs().max()) > 0
    N = 20000
    off = 0
    # per parameter tensor (flat-buffer order: xyz, f_dc, f_rest, opacity, scaling, rotation, then the MLP)
    for name, n in (("xyz", 3 * N), ("f_dc", 3 * N), ("f_rest", 45 * N), ("opacity", N), ("scaling", 3 * N),
                    ("rotation", 4 * N), ("mlp", a.numel() - 59 * N)):
        x, y = a[off:off + n], b[off:off + n]
        off += n
        err = float((x - y).abs().max() / x.abs().max().clamp_min(1e-30))
        assert err < 1e-5, (name, err)
