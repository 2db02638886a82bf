"""Dev: per-warp tile-start timestamps of block (0,0) of the warp-specialised backward kernel."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L
lib = ctypes.CDLL(sys.argv[1])
for name, (res, args) in L.SIGNATURES.items():
    fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
lib.curl_dev_set_debug.argtypes = [ctypes.c_void_p]
B, H, W = 8, 2160, 3840
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1234)
img = torch.rand(B, 3, H, W, generator=g, device=dev)
mask = (torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1)
Lk, Rk, Hk = (0.1 * torch.randn(B, n, generator=g, device=dev) for n in (48, 48, 64))
gout = torch.randn(B, 3, H, W, generator=g, device=dev); greg = torch.ones(B, device=dev)
gimg = torch.empty_like(img); gL = torch.empty_like(Lk); gR = torch.empty_like(Rk); gH = torch.empty_like(Hk)
wsb = lib.curl_bwd_workspace_bytes(B, H, W); ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
dbg = torch.zeros(32 * 64, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
for rep in range(3):
    lib.curl_dev_set_debug(dbg.data_ptr() if rep == 2 else None)
    rc = lib.curl_layer_bwd_f32(img.data_ptr(), mask.data_ptr(), 1, Lk.data_ptr(), 48, Rk.data_ptr(), 48, Hk.data_ptr(), 64,
                                gout.data_ptr(), greg.data_ptr(), gimg.data_ptr(), gL.data_ptr(), gR.data_ptr(), gH.data_ptr(),
                                ws.data_ptr(), wsb, B, H, W, 16, 16, 16, st)
    assert rc == 0
torch.cuda.synchronize()
t = dbg.cpu().view(32, 64)
nw = int((t[:, 0] != 0).sum())
t0 = int(t[:nw, 0].min())
print("warps", nw)
for w in range(nw):
    row = (t[w] - t0).tolist()
    d = [row[i + 1] - row[i] for i in range(40, 63)]
    print("warp %2d smsp %d: start@it40 %7d  iter time mean %6.0f min %5d max %5d | it40..47 starts rel warp0: %s" % (
        w, w % 4, row[40], sum(d) / len(d), min(d), max(d), [row[i] - int(t[0, i] - t0) for i in range(40, 48)]))
