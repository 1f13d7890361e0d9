"""Engineering: steady-state time of one fused-GEMM launch shape, launched back to back over rotating buffer sets
(cold L2 for the inputs, like inside the forward).   python tools/bench_linear_loop.py Cin N [plain|resid|aggr] [sets]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from super_resolution_b200 import _lib
L = _lib.lib()
p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
B, H, W = 16, 64, 64
Cin, N = int(sys.argv[1]), int(sys.argv[2])
mode = sys.argv[3] if len(sys.argv) > 3 else "plain"
nsets = int(sys.argv[4]) if len(sys.argv) > 4 else 6
M = B * H * W
pitch = ((Cin + 63) // 64) * 64
opitch = ((N + 63) // 64) * 64
w = torch.randn(N, Cin, device="cuda") * 0.05
b = torch.randn(N, device="cuda")
g = torch.ones(N, device="cuda"); be = torch.zeros(N, device="cuda")
sc = torch.rand(B, N, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
sets = []
for _ in range(nsets):
    x = torch.zeros(M, pitch, dtype=torch.bfloat16, device="cuda"); x[:, :Cin] = torch.randn(M, Cin, device="cuda").bfloat16()
    d = dict(x=x)
    if mode == "plain":
        d["out"] = torch.zeros(M, opitch, dtype=torch.bfloat16, device="cuda")
    else:
        d["r"] = torch.randn(M, N, device="cuda"); d["o32"] = torch.empty(M, N, device="cuda")
        d["u"] = torch.randn(M, opitch, device="cuda").bfloat16()
        d["lo"] = torch.empty(M, opitch, dtype=torch.bfloat16, device="cuda")
        d["hd"] = torch.empty(M, 16, dtype=torch.bfloat16, device="cuda"); d["cs"] = torch.empty(M // 128, 16, device="cuda")
    sets.append(d)
def run(d):
    if mode == "plain":
        _lib.check(L.hatb200_linear_fused(p(d["x"]), pitch, Cin, B, H, W, p(w), p(b), N, 0, None, None, 0, None, 0.0,
                                          None, p(d["out"]), opitch, 0, None, None, None, 0, None, None, st))
    else:
        eca = mode == "aggr"
        _lib.check(L.hatb200_linear_fused(p(d["x"]), pitch, Cin, B, H, W, p(w), p(b), N, 0, p(d["r"]), p(d["u"]) if eca else None,
                                          opitch, p(sc) if eca else None, 0.01, p(d["o32"]), None, 0, 0, p(g), p(be),
                                          p(d["lo"]), opitch, None if eca else p(d["hd"]), None if eca else p(d["cs"]), st))
for k in range(nsets):
    run(sets[k])
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(60):
    run(sets[k % nsets])
e1.record()
torch.cuda.synchronize()
print(f"{mode} {Cin}->{N}: {e0.elapsed_time(e1) * 1000 / 60:.1f} us per launch back to back over {nsets} buffer sets",
      " ".join(f"{k[8:]}={v}" for k, v in os.environ.items() if k.startswith("HATB200_G3")))
