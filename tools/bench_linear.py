"""Time one fused Linear launch (the C-ABI test entry) at the bench workload's token count, L2 flushed between runs.
    python tools/bench_linear.py Cin N [plain|gelu|resid|aggr] [iters]
plain: bias -> bf16; resid: + fp32 residual -> fp32 + LayerNorm bf16 + 16-channel head (fc2 / proj shape);
aggr: + ECA-scaled CAB branch, no head (ESC aggr shape).  Engineering knobs are read by the library from the
environment (HATB200_G3_OPS / _STGBUFS / _RESIDENT, HATB200_EPI_GENERIC).
"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from super_resolution_b200 import _lib
L = _lib.lib()
p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
B, H, W = 16, 64, 64
Cin, N = int(sys.argv[1]), int(sys.argv[2])
mode = sys.argv[3] if len(sys.argv) > 3 else "plain"
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 20
M = B * H * W
pitch = ((Cin + 63) // 64) * 64
x = torch.zeros(M, pitch, dtype=torch.bfloat16, device="cuda"); x[:, :Cin] = torch.randn(M, Cin, device="cuda").bfloat16()
w = torch.randn(N, Cin, device="cuda") * 0.05
b = torch.randn(N, device="cuda")
opitch = ((N + 63) // 64) * 64
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
if mode in ("plain", "gelu"):
    out = torch.zeros(M, opitch, dtype=torch.bfloat16, device="cuda")
    run = lambda: _lib.check(L.hatb200_linear_fused(p(x), pitch, Cin, B, H, W, p(w), p(b), N, 1 if mode == "gelu" else 0, None, None, 0,
                                                    None, 0.0, None, p(out), opitch, 0, None, None, None, 0, None, None, st))
    nbytes = M * (pitch + opitch) * 2
else:
    r = torch.randn(M, N, device="cuda"); o32 = torch.empty(M, N, device="cuda")
    u = torch.randn(M, opitch, device="cuda").bfloat16(); sc = torch.rand(B, N, device="cuda")
    g = torch.ones(N, device="cuda"); be = torch.zeros(N, device="cuda")
    lo = torch.empty(M, opitch, dtype=torch.bfloat16, device="cuda")
    hd = torch.empty(M, 16, dtype=torch.bfloat16, device="cuda"); cs = torch.empty(M // 128, 16, device="cuda")
    eca = mode == "aggr"
    run = lambda: _lib.check(L.hatb200_linear_fused(p(x), pitch, Cin, B, H, W, p(w), p(b), N, 0, p(r), p(u) if eca else None,
                                                    opitch, p(sc) if eca else None, 0.01, p(o32), None, 0, 0, p(g), p(be),
                                                    p(lo), opitch, None if eca else p(hd), None if eca else p(cs), st))
    nbytes = M * (pitch * 2 + N * 8 + opitch * 2 + (opitch * 2 if eca else 0))
for _ in range(3):
    run()
ts = []
for _ in range(iters):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1000)
ts.sort()
med = ts[len(ts) // 2]
knobs = " ".join(f"{k[8:]}={v}" for k, v in os.environ.items() if k.startswith("HATB200_") and k != "HATB200_TRACE")
print(f"{mode:6s} {Cin:4d}->{N:4d}  median {med:6.1f} us  min {ts[0]:6.1f}  {nbytes / med / 1e3:7.0f} GB/s  [{knobs}]")
