"""CPU: the dopri5 + adjoint solver (product package, torch ops only) on a linear system with a closed-form solution."""
import torch

from conftest import pkg


class _Lin(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Parameter(torch.tensor([[-0.5, 1.0], [-1.0, -0.5]], dtype=torch.float64))

    def forward(self, t, y):
        return y @ self.a.T * (1 + 0 * t)


def test_dopri5_and_adjoint_on_a_linear_system():
    ode = pkg("ode")
    f = _Lin()
    y0 = torch.tensor([[1.0, 0.0], [0.3, -0.7]], dtype=torch.float64, requires_grad=True)
    t = torch.tensor([0.0, 0.5, 1.3], dtype=torch.float64)
    ys = ode.odeint_adjoint(f, y0, t, rtol=1e-9, atol=1e-11)
    exact = torch.stack([y0.detach() @ torch.matrix_exp(f.a.detach() * tt).T for tt in t.tolist()])
    assert float((ys.detach() - exact).abs().max()) < 1e-8
    w = torch.randn(ys.shape, dtype=torch.float64, generator=torch.Generator().manual_seed(0))
    (ys * w).sum().backward()
    f2 = _Lin()
    y02 = y0.detach().clone().requires_grad_(True)
    (torch.stack([y02 @ torch.matrix_exp(f2.a * tt).T for tt in t.tolist()]) * w).sum().backward()
    assert float((f.a.grad - f2.a.grad).abs().max()) < 1e-7
    assert float((y0.grad - y02.grad).abs().max()) < 1e-7


def test_backward_in_time_and_loose_tolerance():
    ode = pkg("ode")
    f = _Lin()
    y0 = torch.tensor([[0.2, 0.9]], dtype=torch.float64)
    t = torch.tensor([1.0, 0.25, 0.0], dtype=torch.float64)
    ys = ode.odeint(f, y0, t, rtol=1e-4, atol=1e-6)
    exact = torch.stack([y0 @ torch.matrix_exp(f.a.detach() * (tt - 1.0)).T for tt in t.tolist()])
    assert float((ys - exact).abs().max()) < 1e-3
