# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""Per-CTA timeline of one persistent GEMM launch (engineering aid, HATB200_GEMM_TRACE=1).
    python tools/trace_gemm.py Cin N [plain|resid|aggr]
Slots (us since the first event): 0/1 MMA warp start / weights resident; 8+2i / 9+2i MMA of local tile i issued /
committed; 24+16g+4k+{0,1,3,2}: epilogue warpgroup g, its k-th tile: waiting, accumulator ready, pass 1 done, tile done.
"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
NO_TRACE = bool(os.environ.get("HATB200_NO_TRACE"))      # plain run of the same launch (ncu target)
if not NO_TRACE:
    os.environ["HATB200_GEMM_TRACE"] = "1"
import numpy as np, torch
from super_resolution_b200 import _lib
lib = C.CDLL(_lib.LIB_PATH)
L = _lib.lib()
p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
B, H, W = 16, 64, 64
Cin, N = int(sys.argv[1]) if len(sys.argv) > 1 else 180, int(sys.argv[2]) if len(sys.argv) > 2 else 720
mode = sys.argv[3] if len(sys.argv) > 3 else "plain"
M = B * H * W
pitch = ((Cin + 63) // 64) * 64
x = torch.zeros(M, pitch, dtype=torch.bfloat16, device="cuda"); x[:, :Cin] = torch.randn(M, Cin, device="cuda").bfloat16()
w = torch.randn(N, Cin, device="cuda") * 0.05
b = torch.randn(N, device="cuda")
opitch = ((N + 63) // 64) * 64
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
if mode == "plain":
    out = torch.zeros(M, opitch, dtype=torch.bfloat16, device="cuda")
    run = lambda: _lib.check(L.hatb200_linear_fused(p(x), pitch, Cin, B, H, W, p(w), p(b), N, 0, None, None, 0, None, 0.0,
                                                    None, p(out), opitch, 0, None, None, None, 0, None, None, st))
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
for _ in range(3):
    flush.zero_()
    run()
torch.cuda.synchronize()
if NO_TRACE:
    print("ran", mode, Cin, N)
    sys.exit(0)
buf = np.zeros((1024, 64), dtype=np.uint64)
lib.hatb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_void_p]
rc = lib.hatb200_debug_gemm_trace(buf.ctypes.data, None)
t = buf.astype(np.int64)
live = t[:, 0] > 0
t0 = t[live, 0].min()
knobs = " ".join(f"{k[8:]}={v}" for k, v in os.environ.items() if k.startswith("HATB200_G3") or k.startswith("HATB200_EPI"))
print(mode, Cin, N, "ctas", live.sum(), "span us", (t[live, :56].max() - t0) / 1000, "[" + knobs + "]")
if not os.environ.get("VERBOSE"):
    sys.exit(0)
for cta in list(np.nonzero(live)[0][:2]) + [int(np.nonzero(live)[0][-1])]:
    r_ = t[cta]
    ev = {k: int(r_[k] - t0) for k in range(56) if r_[k] > 0}
    print("cta", cta, {k: round(v / 1000, 2) for k, v in sorted(ev.items())})
    for g in range(2):   # epilogue warpgroup totals over its tiles, cycles of SM clock -> us at 1.965 GHz
        ring, stg, p1, p2 = (float(r_[56 + 4 * g + j]) / 1965.0 for j in range(4))
        print("   wg%d: pass1 %.2f us (ring wait %.2f, staging wait %.2f)  pass2 %.2f us" % (g, p1, ring, stg, p2))
