"""ORACLE — TEST INFRASTRUCTURE ONLY.  numpy restatement of the reference deformation networks.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product package never does.

Follows /root/reference:
  * Embedder.embed            utils/time_utils.py:24-54   (include_input, log-sampled 2^k, [sin, cos] per freq)
  * DeformNetworkBaseline     utils/time_utils.py:56-127  (timenet :74-76, trunk :108-112, heads :123-125)
  * DeformNetwork             utils/time_utils.py:129-201 (returns d_xyz, 0, 0)
  * 6-DoF branch              utils/time_utils.py:114-121 (`w / theta + 1e-5` precedence kept as written)
  * exp_so3 / exp_se3         utils/rigid_utils.py:42-83
Parameters are a dict with the reference's state_dict keys (`timenet.0.weight`, `linear.3.bias`, ...).

PINNED: tests/test_oracle_mlp.py checks forward outputs and every parameter gradient of this file
against golden vectors produced by importing the reference's own torch module in the build container
(tests/golden/make_golden.py -> tests/golden/mlp_*.npz).
"""
import numpy as np


def embed(x, L):
    """[x, sin(2^0 x), cos(2^0 x), ..., sin(2^(L-1) x), cos(2^(L-1) x)] along the last axis."""
    outs = [x]
    for k in range(L):
        f = x.dtype.type(2.0 ** k)
        outs.append(np.sin(x * f))
        outs.append(np.cos(x * f))
    return np.concatenate(outs, axis=-1)


def embed_grad(x, g, L):
    """d(sum(g*embed(x)))/dx."""
    d = x.shape[-1]
    out = g[..., :d].copy()
    o = d
    for k in range(L):
        f = x.dtype.type(2.0 ** k)
        out += g[..., o:o + d] * np.cos(x * f) * f
        o += d
        out -= g[..., o:o + d] * np.sin(x * f) * f
        o += d
    return out


def _skew(w):
    z = np.zeros(w.shape[0], w.dtype)
    return np.stack([z, -w[:, 2], w[:, 1], w[:, 2], z, -w[:, 0], -w[:, 1], w[:, 0], z], -1).reshape(-1, 3, 3)


def exp_se3(S, theta):
    """utils/rigid_utils.py:60-83.  S [N,6], theta [N,1] -> [N,4,4]."""
    w, v = S[:, :3], S[:, 3:]
    W = _skew(w)
    W2 = W @ W
    I = np.eye(3, dtype=S.dtype)[None]
    th = theta.reshape(-1, 1, 1)
    R = I + np.sin(th) * W + (1.0 - np.cos(th)) * W2
    p = (th * I + (1.0 - np.cos(th)) * W + (th - np.sin(th)) * W2) @ v[:, :, None]
    T = np.zeros((S.shape[0], 4, 4), S.dtype)
    T[:, :3, :3] = R
    T[:, :3, 3:] = p
    T[:, 3, 3] = 1.0
    return T


class DeformRef:
    """Forward + analytic backward of the reference network in numpy."""

    def __init__(self, params, is_blender=True, is_6dof=False, baseline=True, dtype=np.float64):
        self.p = {k: np.asarray(v, dtype=dtype) for k, v in params.items()}
        self.is_blender, self.is_6dof, self.baseline, self.dtype = is_blender, is_6dof, baseline, dtype
        self.tL = 6 if is_blender else 10

    def forward(self, x, t):
        p, dt = self.p, self.dtype
        x = np.asarray(x, dt)
        t = np.asarray(t, dt).reshape(-1, 1)
        if t.shape[0] == 1 and x.shape[0] != 1:
            t = np.broadcast_to(t, (x.shape[0], 1)).copy()
        c = {"x": x, "t": t}
        te = embed(t, self.tL)
        c["te"] = te
        if self.is_blender:
            z = te @ p["timenet.0.weight"].T + p["timenet.0.bias"]
            c["tn_h"] = np.maximum(z, 0)
            tf = c["tn_h"] @ p["timenet.2.weight"].T + p["timenet.2.bias"]
        else:
            tf = te
        xe = embed(x, 10)
        pe = np.concatenate([xe, tf], -1)
        c["pe"] = pe
        h = pe
        c["in"], c["act"], c["z"] = [], [], []
        for i in range(8):
            c["in"].append(h)
            z = h @ p["linear.%d.weight" % i].T + p["linear.%d.bias" % i]
            a = np.maximum(z, 0)
            c["z"].append(z)
            c["act"].append(a)
            h = a
            if i == 4:
                h = np.concatenate([pe, a], -1)
        c["h"] = h
        if self.is_6dof:
            w = h @ p["branch_w.weight"].T + p["branch_w.bias"]
            v = h @ p["branch_v.weight"].T + p["branch_v.bias"]
            theta = np.sqrt((w * w).sum(-1, keepdims=True))
            c["w_raw"], c["v_raw"], c["theta"] = w, v, theta
            S = np.concatenate([w / theta + dt(1e-5), v / theta + dt(1e-5)], -1)
            d_xyz = exp_se3(S, theta)
        else:
            d_xyz = h @ p["gaussian_warp.weight"].T + p["gaussian_warp.bias"]
        self.cache = c
        if self.baseline:
            rot = h @ p["gaussian_rotation.weight"].T + p["gaussian_rotation.bias"]
            sc = h @ p["gaussian_scaling.weight"].T + p["gaussian_scaling.bias"]
            return d_xyz, rot, sc
        return d_xyz, 0, 0

    def _se3_backward(self, g):
        """g [N,4,4] -> gradients w.r.t. raw head outputs w, v (finite-difference-checked in tests)."""
        c, dt = self.cache, self.dtype
        w_raw, v_raw, theta = c["w_raw"], c["v_raw"], c["theta"]
        wn = w_raw / theta + dt(1e-5)
        vn = v_raw / theta + dt(1e-5)
        W = _skew(wn)
        W2 = W @ W
        th = theta.reshape(-1, 1, 1)
        s, co = np.sin(th), np.cos(th)
        gR, gp = g[:, :3, :3], g[:, :3, 3]
        I = np.eye(3, dtype=dt)[None]
        P = th * I + (1 - co) * W + (th - s) * W2
        g_vn = (P.transpose(0, 2, 1) @ gp[:, :, None])[:, :, 0]
        gP = gp[:, :, None] * vn[:, None, :]
        gW = s * gR + (1 - co) * gP
        gW2 = (1 - co) * gR + (th - s) * gP
        gW = gW + gW2 @ W.transpose(0, 2, 1) + W.transpose(0, 2, 1) @ gW2
        g_th = ((co * W + s * W2) * gR).sum((1, 2)) + ((I + s * W + (1 - co) * W2) * gP).sum((1, 2))
        g_wn = np.stack([gW[:, 2, 1] - gW[:, 1, 2], gW[:, 0, 2] - gW[:, 2, 0], gW[:, 1, 0] - gW[:, 0, 1]], -1)
        # wn = w/theta + eps, vn = v/theta + eps, theta = |w|
        g_w = g_wn / theta
        g_v = g_vn / theta
        g_theta = g_th[:, None] - (g_wn * w_raw).sum(-1, keepdims=True) / theta ** 2 \
            - (g_vn * v_raw).sum(-1, keepdims=True) / theta ** 2
        g_w = g_w + g_theta * w_raw / theta
        return g_w, g_v

    def backward(self, g_dxyz, g_rot=None, g_scale=None, want_input_grads=False, masks=None):
        """Returns a dict of parameter gradients (reference state_dict keys); optionally dL/dx, dL/dt.
        `masks` ([8,N,256] bool) overrides the ReLU activation pattern (z > 0) of the trunk, so that a
        reduced-precision forward can be checked for arithmetic accuracy separately from the (inherent)
        sign flips of pre-activations that sit within rounding noise of zero."""
        p, c, dt = self.p, self.cache, self.dtype
        grads = {}
        h = c["h"]
        gh = np.zeros_like(h)

        def lin_bwd(name, inp, gout):
            grads[name + ".weight"] = gout.T @ inp
            grads[name + ".bias"] = gout.sum(0)
            return gout @ p[name + ".weight"]

        if self.is_6dof:
            g_w, g_v = self._se3_backward(np.asarray(g_dxyz, dt))
            gh += lin_bwd("branch_w", h, g_w)
            gh += lin_bwd("branch_v", h, g_v)
        else:
            gh += lin_bwd("gaussian_warp", h, np.asarray(g_dxyz, dt))
        if self.baseline:
            if g_rot is not None:
                gh += lin_bwd("gaussian_rotation", h, np.asarray(g_rot, dt))
            if g_scale is not None:
                gh += lin_bwd("gaussian_scaling", h, np.asarray(g_scale, dt))
        g_pe = np.zeros_like(c["pe"])
        g = gh
        for i in range(7, -1, -1):
            gz = g * ((c["act"][i] > 0) if masks is None else masks[i])
            g = lin_bwd("linear.%d" % i, c["in"][i], gz)
            if i == 5:
                npe = c["pe"].shape[1]
                g_pe += g[:, :npe]
                g = g[:, npe:]
        g_pe += g
        g_tf = g_pe[:, 63:]
        if self.is_blender:
            g_hid = lin_bwd("timenet.2", c["tn_h"], g_tf)
            g_hid = g_hid * (c["tn_h"] > 0)
            g_te = lin_bwd("timenet.0", c["te"], g_hid)
        else:
            g_te = g_tf
        if want_input_grads:
            grads["_x"] = embed_grad(c["x"], g_pe[:, :63], 10)
            grads["_t"] = embed_grad(c["t"], g_te, self.tL)
        return grads


def param_shapes(is_blender=True, is_6dof=False):
    """state_dict keys and shapes of the reference module (verified against the real module by
    tests/golden/make_golden.py)."""
    pe = 63 + (30 if is_blender else 21)
    s = {}
    if is_blender:
        s["timenet.0.weight"], s["timenet.0.bias"] = (256, 13), (256,)
        s["timenet.2.weight"], s["timenet.2.bias"] = (30, 256), (30,)
    for i in range(8):
        s["linear.%d.weight" % i] = (256, pe if i == 0 else (256 + pe if i == 5 else 256))
        s["linear.%d.bias" % i] = (256,)
    if is_6dof:
        s["branch_w.weight"], s["branch_w.bias"] = (3, 256), (3,)
        s["branch_v.weight"], s["branch_v.bias"] = (3, 256), (3,)
    else:
        s["gaussian_warp.weight"], s["gaussian_warp.bias"] = (3, 256), (3,)
    s["gaussian_rotation.weight"], s["gaussian_rotation.bias"] = (4, 256), (4,)
    s["gaussian_scaling.weight"], s["gaussian_scaling.bias"] = (3, 256), (3,)
    return s


def seeded_params(seed, is_blender=True, is_6dof=False, head_scale=1.0):
    """Deterministic, platform-independent parameters: uniform(-1/sqrt(fan_in), 1/sqrt(fan_in)) like
    nn.Linear's default scale, drawn from numpy's PCG64 (bit-reproducible everywhere)."""
    rng = np.random.default_rng(seed)
    out = {}
    for k, shp in param_shapes(is_blender, is_6dof).items():
        layer = k.rsplit(".", 1)[0]
        fan_in = param_shapes(is_blender, is_6dof)[layer + ".weight"][1]
        bound = 1.0 / np.sqrt(fan_in)
        v = rng.uniform(-bound, bound, size=shp).astype(np.float32)
        if layer in ("branch_w", "branch_v", "gaussian_rotation", "gaussian_scaling"):
            v = (v * head_scale).astype(np.float32)
        out[k] = v
    return out
