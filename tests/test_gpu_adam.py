"""GPU: fused multi-tensor Adam (csrc/adam.cu through the C ABI) against torch.optim.Adam — the optimizer the
reference constructs (scene/gaussian_model.py:136, scene/deform_model.py:295,337) — on the same parameters,
gradients, per-group learning rates and the reference's eps."""
import pytest
import torch

from conftest import pkg


def _groups(dev, seed):
    g = torch.Generator().manual_seed(seed)
    shapes = {"xyz": (5001, 3), "f_dc": (5001, 1, 3), "f_rest": (5001, 15, 3), "opacity": (5001, 1),
              "scaling": (5001, 3), "rotation": (5001, 4), "w": (256, 93), "b": (256,), "odd": (7,)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3,
           "w": 8e-4, "b": 8e-4, "odd": 1e-2}
    return [{"params": [torch.randn(*s, generator=g).to(dev).requires_grad_(True)], "lr": lrs[n], "name": n}
            for n, s in shapes.items()]


@pytest.mark.gpu
@pytest.mark.parametrize("weight_decay", [0.0, 1e-2])
def test_fused_adam_matches_torch(weight_decay):
    optim = pkg("optim")
    dev = torch.device("cuda", 0)
    ga, gb = _groups(dev, 0), _groups(dev, 0)
    ref = torch.optim.Adam(ga, lr=0.0, eps=1e-15, weight_decay=weight_decay)
    ours = optim.Adam(gb, lr=0.0, eps=1e-15, weight_decay=weight_decay)
    gen = torch.Generator().manual_seed(1)
    for it in range(5):
        for A, B in zip(ga, gb):
            grad = (torch.randn(A["params"][0].shape, generator=gen) * 10.0 ** (-(it % 3) * 2)).to(dev)
            if it == 2 and A["name"] == "opacity":
                A["params"][0].grad = B["params"][0].grad = None     # a parameter without gradient is skipped
                continue
            A["params"][0].grad, B["params"][0].grad = grad.clone(), grad.clone()
            A["lr"] = B["lr"] = A["lr"] * 0.97                        # the schedulers rewrite group['lr'] every step
        ref.step()
        ours.step()
        for A, B in zip(ga, gb):
            pa, pb = A["params"][0], B["params"][0]
            assert (pa - pb).abs().max() <= 2e-6 * pa.abs().max().clamp_min(1.0), (it, A["name"])
            sa, sb = ref.state[pa], ours.state[pb]
            assert float(sa["step"]) == float(sb["step"])
            for k in ("exp_avg", "exp_avg_sq"):   # max-norm relative: elements that cancel to ~0 carry no relative meaning
                assert (sa[k] - sb[k]).abs().max() <= 2e-6 * sa[k].abs().max(), (it, A["name"], k)
    # state_dict interchange with torch.optim.Adam (checkpoints, densification surgery)
    fresh = torch.optim.Adam(_groups(dev, 0), lr=0.0, eps=1e-15, weight_decay=weight_decay)
    fresh.load_state_dict(ours.state_dict())
    with pytest.raises(RuntimeError):
        cpu = optim.Adam([torch.zeros(3, requires_grad=True)], lr=1e-3)
        cpu.param_groups[0]["params"][0].grad = torch.ones(3)
        cpu.step()
