"""Clock stamps of one tile of the fused MLP forward (CTA 0, its second tile): D3G_FWD_DBG=32."""
import ctypes as C, importlib, os, sys
import torch
os.environ["D3G_FWD_DBG"] = str(32 | int(os.environ.get("EXTRA", "0")))
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
_lib = importlib.import_module("deformable-3d-gaussians_b200._lib")
lib = _lib.load()
save = int(sys.argv[1]) if len(sys.argv) > 1 else 0
N = 100000
net = deform.DeformNetworkBaseline(is_blender=True).cuda()
x = torch.rand(N, 3, device="cuda") * 2.6 - 1.3
t = torch.tensor([0.37], device="cuda")
cfg = net._cfg()
packed = net._packed(x.device)
wsb = C.c_size_t()
_lib.check(lib.d3g_mlp_workspace_bytes_times(C.byref(cfg), N, 1, save, C.byref(wsb)))
ws = torch.zeros(wsb.value, dtype=torch.uint8, device="cuda")
ps = deform._fill_ptr_struct(_lib.MlpParams(), deform._param_list(net))
o1, o2, o3 = torch.empty(N, 3, device="cuda"), torch.empty(N, 4, device="cuda"), torch.empty(N, 3, device="cuda")
for _ in range(3):
    _lib.check(lib.d3g_mlp_forward_times(C.byref(cfg), _lib.ptr(packed), C.byref(ps), C.c_int64(N), 1, _lib.ptr(x), _lib.ptr(t),
                                         _lib.ptr(ws), save, _lib.ptr(o1), _lib.ptr(o2), _lib.ptr(o3), C.c_void_p(0)))
torch.cuda.synchronize()
tl = ws[64:64 + 224 * 4].view(torch.int32).cpu().tolist()
print("MMA thread: tile start %d | commit after layer l:" % tl[0], tl[1:9])
print("layer 1 MMA: a_ready[kc] seen:", tl[16:24])
print("layer 1 MMA: weights ready   :", tl[24:32])
for l in range(3):
    for cg in range(2):
        b = 40 + l * 24 + cg * 12
        print("epilogue layer %d warp cg=%d: acc_full seen %d | per chunk (after ld, after arrive):" % (l, cg, tl[b]), tl[b + 1:b + 9])
