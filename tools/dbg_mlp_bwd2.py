"""Decode the saved wgrad images (h, dz, mask) of tile 0 and compare with the oracle's intermediates."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mlp_ref
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")

def al(x, a=1024): return (x + a - 1) // a * a

def ws_offsets(rows):
    tiles = max((rows + 127) // 128, 1)
    off = {}; o = 0
    def take(name, b):
        nonlocal o
        off[name] = o; o += al(b)
    take("status", 1024); take("tfeat", rows * 128); take("thidden", rows * 1024)
    take("pe", tiles * 49152); take("h", 8 * tiles * 131072); take("mask", 8 * tiles * 8 * 128 * 4)
    take("heads_out", tiles * 128 * 64); take("dz", 8 * tiles * 131072); take("dout", tiles * 16384)
    return off, tiles

def decode(img, C):
    """img: float32 array of one tile image -> [128, C]"""
    out = np.zeros((128, C), np.float32)
    p = np.arange(128)[:, None]; f = np.arange(C)[None, :]
    byte = (p // 32) * (C // 32) * 4096 + (f // 32) * 4096 + (p % 32) * 128 + ((((f % 32) // 8) ^ (p & 3)) << 5) + (f % 8) * 4
    return img[byte // 4]

N = 128
params = mlp_ref.seeded_params(5)
net = deform.DeformNetworkBaseline(is_blender=True)
net.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
net.debug = True
net = net.cuda()
rng = np.random.default_rng(1)
x = rng.uniform(-1.3, 1.3, size=(N, 3)).astype(np.float32)
gd = rng.standard_normal((N, 3)).astype(np.float32)
gr, gs = (rng.standard_normal((N, k)).astype(np.float32) for k in (4, 3))
out = net(torch.tensor(x).cuda(), torch.full((1, 1), 0.5).cuda().expand(N, 1))
((out[0] * torch.tensor(gd).cuda()).sum() + (out[1] * torch.tensor(gr).cuda()).sum() + (out[2] * torch.tensor(gs).cuda()).sum()).backward()
ws = deform._FusedDeform.last_ws.cpu().numpy()
off, tiles = ws_offsets(N)
ref = mlp_ref.DeformRef(params, True, False, True, dtype=np.float64)
ref.forward(x, np.array([[0.5]]))
c = ref.cache
# recompute oracle dz per layer
p = ref.p
gh = gd @ p["gaussian_warp.weight"] + gr @ p["gaussian_rotation.weight"] + gs @ p["gaussian_scaling.weight"]
g = gh
for l in range(7, -1, -1):
    gz = g * (c["act"][l] > 0)
    himg = ws[off["h"] + l * tiles * 131072: off["h"] + l * tiles * 131072 + 131072].view(np.float32)
    dzimg = ws[off["dz"] + l * tiles * 131072: off["dz"] + l * tiles * 131072 + 131072].view(np.float32)
    h_got, dz_got = decode(himg, 256), decode(dzimg, 256)
    eh = np.abs(h_got - c["act"][l]).max() / np.abs(c["act"][l]).max()
    ed = np.abs(dz_got - gz).max() / np.abs(gz).max()
    # where are the errors
    bad = np.argwhere(np.abs(dz_got - gz) > 5e-3 * np.abs(gz).max())
    print("layer %d: h rel %.2e  dz rel %.2e  bad count %d  first %s" % (l, eh, ed, len(bad), bad[:6].tolist()))
    if len(bad):
        i, j = bad[0]
        print("    got %.5f ref %.5f  act %.5f" % (dz_got[i, j], gz[i, j], c["act"][l][i, j]))
        print("    bad cols hist (chunk):", np.bincount(bad[:, 1] // 32, minlength=8).tolist(), " rows hist (quarter):", np.bincount(bad[:, 0] // 32, minlength=4).tolist())
    g = gz @ p["linear.%d.weight" % l]
    if l == 5:
        g = g[:, 93:]
doutimg = ws[off["dout"]: off["dout"] + 16384].view(np.float32)
print("dout:", np.abs(decode(doutimg, 32)[:, :10] - np.concatenate([gd, gr, gs], 1)).max())
