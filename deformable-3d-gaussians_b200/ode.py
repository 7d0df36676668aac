"""Adaptive Dormand-Prince (dopri5) initial-value solver with the continuous adjoint, for the `ode_batch*`
configurations: scene/deform_model.py:183-207 calls `torchdiffeq.odeint_adjoint(self.deform, xyz, t_interval,
rtol=, atol=)` (torchdiffeq is a third-party dependency that is not part of the reference tree and is not installed
here; DeformModelODE uses it when importable).  This module provides the same call form —
`odeint_adjoint(func, y0, t, rtol=, atol=)` -> [len(t), *y0.shape], differentiable w.r.t. y0, t and func's parameters
— following the published algorithm of that library's default method: dopri5 with the embedded 4th-order error
estimate, RMS error norm against atol + rtol * max(|y0|, |y1|), step-size controller with safety 0.9 and growth
limits [0.2, 10], Hairer's initial-step heuristic, a 4th-order interpolant for the requested output times, and the
adjoint method of Chen et al. 2018 (the augmented state (y, a, dL/dt, dL/dtheta) integrated backwards between output
times).  Every right-hand-side evaluation is one launch of the fused MLP (DeformNetworkODE.forward), every adjoint
evaluation one fused forward + backward with input gradients.  Solver parity with torchdiffeq: unpinned (no copy of
it to run here); tests check the solver against a fixed-step fp64 Runge-Kutta integration of the numpy oracle and
the adjoint against finite differences."""
import torch

# Dormand-Prince 5(4) tableau
_ALPHA = [1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
_BETA = [[1 / 5],
         [3 / 40, 9 / 40],
         [44 / 45, -56 / 15, 32 / 9],
         [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
         [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
         [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84]]
_C_SOL = [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0]
_C_ERR = [35 / 384 - 1951 / 21600, 0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720,
          -2187 / 6784 - -12231 / 42400, 11 / 84 - 649 / 6300, -1.0 / 60.0]
_C_MID = [6025192743 / 30085553152 / 2, 0, 51252292925 / 65400821598 / 2, -2691868925 / 45128329728 / 2,
          187940372067 / 1594534317056 / 2, -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2]


def _rms(x):
    return x.pow(2).mean().sqrt() if x.numel() else x.new_zeros(())


class _State:
    """A tuple of tensors treated as one vector; the error norm is the largest RMS norm over its first `n_norm`
    members (all of them by default)."""

    def __init__(self, parts, n_norm=None):
        self.parts = tuple(parts)
        self.n_norm = len(self.parts) if n_norm is None else n_norm

    def axpy(self, coeffs, ks, dt):
        out = []
        for i, p in enumerate(self.parts):
            acc = None
            for c, k in zip(coeffs, ks):
                if c == 0:
                    continue
                term = k.parts[i] * (c * dt)
                acc = term if acc is None else acc + term
            out.append(p if acc is None else p + acc)
        return _State(out, self.n_norm)

    def norm_ratio(self, err, other, rtol, atol):
        r = None
        for e, a, b in list(zip(err.parts, self.parts, other.parts))[:self.n_norm]:
            tol = atol + rtol * torch.max(a.abs(), b.abs())
            v = _rms(e / tol)
            r = v if r is None else torch.max(r, v)
        return r


def _combine(coeffs, ks, dt):
    out = []
    for i in range(len(ks[0].parts)):
        acc = None
        for c, k in zip(coeffs, ks):
            if c == 0:
                continue
            term = k.parts[i] * (c * dt)
            acc = term if acc is None else acc + term
        out.append(acc)
    return _State(out)


def _initial_step(f, t0, y0, f0, rtol, atol, direction):
    """Hairer, Norsett & Wanner, II.4 (order 4 error estimator)."""
    d0 = d1 = None
    n = y0.n_norm
    for y, k in list(zip(y0.parts, f0.parts))[:n]:
        scale = atol + y.abs() * rtol
        a, b = _rms(y / scale), _rms(k / scale)
        d0 = a if d0 is None else torch.max(d0, a)
        d1 = b if d1 is None else torch.max(d1, b)
    h0 = 1e-6 if (float(d0) < 1e-5 or float(d1) < 1e-5) else 0.01 * float(d0) / float(d1)
    y1 = y0.axpy([1.0], [f0], h0 * direction)
    f1 = f(t0 + h0 * direction, y1)
    d2 = None
    for y, k0, k1 in list(zip(y0.parts, f0.parts, f1.parts))[:n]:
        scale = atol + y.abs() * rtol
        v = _rms((k1 - k0) / scale) / h0
        d2 = v if d2 is None else torch.max(d2, v)
    d1f, d2f = float(d1), float(d2)
    h1 = max(1e-6, h0 * 1e-3) if (d1f <= 1e-15 and d2f <= 1e-15) else (0.01 / max(d1f, d2f)) ** (1.0 / 5.0)
    return min(100 * h0, h1)


ADJOINT_FULL_NORM = False
STATS = {"steps": 0, "rejected": 0}   # running counters (diagnostics / tests)


def _solve(f, y0, times, rtol, atol):
    """Integrate the _State y0 through the monotone python-float list `times`; returns the state at every time.
    f(t: float, y: _State) -> _State."""
    direction = 1.0 if times[-1] >= times[0] else -1.0
    out = [y0]
    t, y = times[0], y0
    k1 = f(t, y)
    h = _initial_step(f, t, y, k1, rtol, atol, direction)
    nxt = 1
    steps = 0
    while nxt < len(times):
        if steps > 20000:
            raise RuntimeError("odeint: more than 20000 steps between two output times (tolerances below the "
                               "right-hand side's fp32 noise floor?)")
        if h < 1e-10 * max(1.0, abs(t)):
            raise RuntimeError("odeint: step size underflow at t=%g (rtol/atol below the right-hand side's noise "
                               "floor?)" % t)
        steps += 1
        STATS["steps"] += 1
        dt = h * direction
        ks = [k1]
        for a, row in zip(_ALPHA, _BETA):
            ks.append(f(t + a * dt, y.axpy(row, ks, dt)))
        y1 = y.axpy(_C_SOL, ks, dt)
        err = _combine(_C_ERR, ks, dt)
        ratio = float(y.norm_ratio(err, y1, rtol, atol))
        if ratio <= 1.0:
            t1 = t + dt
            # 4th-order interpolant through (y, k1), the mid-point value and (y1, k7) for output times inside the step
            while nxt < len(times) and (times[nxt] - t1) * direction <= 1e-12 * max(1.0, abs(t1)):
                tq = times[nxt]
                if abs(tq - t1) <= 1e-12 * max(1.0, abs(t1)):
                    out.append(y1)
                else:
                    ymid = y.axpy(_C_MID, ks, dt)
                    x = (tq - t) / dt
                    parts = []
                    for y0p, y1p, ymp, f0p, f1p in zip(y.parts, y1.parts, ymid.parts, ks[0].parts, ks[-1].parts):
                        a = 2 * dt * (f1p - f0p) - 8 * (y1p + y0p) + 16 * ymp
                        b = dt * (5 * f0p - 3 * f1p) + 18 * y0p + 14 * y1p - 32 * ymp
                        c = dt * (f1p - 4 * f0p) - 11 * y0p - 5 * y1p + 16 * ymp
                        d = dt * f0p
                        parts.append((((a * x + b) * x + c) * x + d) * x + y0p)
                    out.append(_State(parts, y.n_norm))
                nxt += 1
            t, y, k1 = t1, y1, ks[-1]
        else:
            STATS["rejected"] += 1
        # controller
        if ratio == 0.0:
            factor = 10.0
        else:
            dfactor = 1.0 if ratio < 1.0 else 0.2
            factor = min(10.0, max(0.9 / ratio ** 0.2, dfactor))
        h = h * factor
        if nxt < len(times):
            # never step past the last requested time by more than needed for the interpolant: clip to the end
            remaining = abs(times[-1] - t)
            if h > remaining and remaining > 0:
                h = remaining
    return out


def odeint(func, y0, t, rtol=1e-7, atol=1e-9):
    """Forward solve only (no graph is recorded).  func(t: 0-dim tensor, y) -> dy/dt."""
    times = [float(v) for v in t.detach().cpu().tolist()]
    with torch.no_grad():
        f = lambda tt, st: _State([func(y0.new_tensor(tt), st.parts[0])])
        sol = _solve(f, _State([y0.detach()]), times, rtol, atol)
    return torch.stack([s.parts[0] for s in sol], 0)


class _Adjoint(torch.autograd.Function):
    @staticmethod
    def forward(ctx, func, rtol, atol, n_params, y0, t, *params):
        ctx.func, ctx.rtol, ctx.atol = func, rtol, atol
        ys = odeint(func, y0, t, rtol, atol)
        ctx.save_for_backward(t, ys, *params)
        return ys

    @staticmethod
    def backward(ctx, grad_ys):
        func, rtol, atol = ctx.func, ctx.rtol, ctx.atol
        # step-size control of the backward solve looks at (y, a) only — the "seminorm" of Kidger et al. 2021, an
        # option of torchdiffeq (adjoint_options=dict(norm="seminorm")): the parameter-gradient integrand is a pure
        # quadrature, and with 522k components evaluated in fp16-operand arithmetic its rounding noise would otherwise
        # drive the controller (measured: 89 instead of ~10 backward steps at rtol 1e-3).  ADJOINT_FULL_NORM = True
        # restores torchdiffeq's default (every component).
        n_norm = None if ADJOINT_FULL_NORM else 2
        t, ys, *params = ctx.saved_tensors
        params = tuple(params)
        times = [float(v) for v in t.detach().cpu().tolist()]
        T = len(times)

        def aug(tt, st):
            # st = (y, a, dL/dt accumulator, dL/dtheta accumulators...)
            y, a = st.parts[0], st.parts[1]
            with torch.enable_grad():
                t_ = y.new_tensor(tt).requires_grad_(True)
                y_ = y.detach().requires_grad_(True)
                fv = func(t_, y_)
                vj = torch.autograd.grad(fv, (t_, y_) + params, -a, allow_unused=True)
            vj = [torch.zeros_like(x) if g is None else g for g, x in zip(vj, (t_, y_) + params)]
            return _State([fv.detach(), vj[1], vj[0].reshape(1)] + [g for g in vj[2:]])

        with torch.no_grad():
            adj_y = grad_ys[T - 1].clone()
            adj_t = [None] * T
            adj_params = [torch.zeros_like(p) for p in params]
            for i in range(T - 1, 0, -1):
                # dL/dt_i = a(t_i) . f(t_i, y_i)
                fi = func(ys.new_tensor(times[i]), ys[i])
                dLdt_i = (fi * adj_y).sum().reshape(1)
                adj_t[i] = dLdt_i
                st = _State([ys[i], adj_y, -dLdt_i] + adj_params, n_norm)
                sol = _solve(aug, st, [times[i], times[i - 1]], rtol, atol)[-1]
                adj_y = sol.parts[1] + grad_ys[i - 1]
                adj_t_prev = sol.parts[2]
                adj_params = list(sol.parts[3:])
                adj_t[i - 1] = adj_t_prev if i - 1 == 0 else None
            grad_t = torch.cat([a if a is not None else ys.new_zeros(1) for a in adj_t]).to(t.dtype)
        return (None, None, None, None, adj_y, grad_t) + tuple(adj_params)


def odeint_adjoint(func, y0, t, rtol=1e-7, atol=1e-9, method=None, options=None, adjoint_params=None, **_unused):
    """`torchdiffeq.odeint_adjoint` call form for the default method (dopri5).  Returns [len(t), *y0.shape]."""
    if method not in (None, "dopri5"):
        raise NotImplementedError("odeint_adjoint: only the default method (dopri5) is implemented")
    if t.dim() != 1 or t.numel() < 2:
        raise ValueError("odeint_adjoint: t must be a 1-D tensor with at least two time points")
    params = tuple(p for p in (adjoint_params if adjoint_params is not None else func.parameters()) if p.requires_grad)
    return _Adjoint.apply(func, float(rtol), float(atol), len(params), y0, t, *params)
