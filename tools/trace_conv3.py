# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""Per-CTA timeline of one persistent 3x3 conv launch (engineering aid, HATB200_GEMM_TRACE=1).
    python tools/trace_conv3.py Cin N
Slots (us since the first event): MMA thread, local tile i: 4i waiting for a free accumulator, 4i+1 waiting for the
halo, 4i+2 first MMA, 4i+3 last MMA issued; epilogue warpgroup of tile i: 16+4i waiting, 17+4i accumulator ready, 18+4i done.
"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HATB200_GEMM_TRACE"] = "1"
import numpy as np, torch
from super_resolution_b200 import _lib
lib = C.CDLL(_lib.LIB_PATH)
L = _lib.lib()
p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
B, H, W = int(os.environ.get("TRACE_B", "16")), 64, 64
Cin, N = int(sys.argv[1]) if len(sys.argv) > 1 else 180, int(sys.argv[2]) if len(sys.argv) > 2 else 60
ACT = int(sys.argv[3]) if len(sys.argv) > 3 else 0          # 1: GELU (CAB conv1)
M = B * H * W
pitch = ((Cin + 63) // 64) * 64
x = torch.zeros(M, pitch, dtype=torch.bfloat16, device="cuda"); x[:, :Cin] = torch.randn(M, Cin, device="cuda").bfloat16()
w = torch.randn(N, Cin, 3, 3, device="cuda") * 0.05
b = torch.randn(N, device="cuda")
opitch = ((N + 63) // 64) * 64
out = torch.zeros(M, opitch, dtype=torch.bfloat16, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.zero_()
    _lib.check(L.hatb200_conv_fused(p(x), pitch, Cin, B, H, W, p(w), p(b), N, 3, ACT, None, None, p(out), opitch, 0, None, None,
                                    None, 0, None, None, None, st))
torch.cuda.synchronize()
buf = np.zeros((1024, 64), dtype=np.uint64)
lib.hatb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_void_p]
lib.hatb200_debug_gemm_trace(buf.ctypes.data, None)
t = buf.astype(np.int64)
live = t[:, 0] > 0
clk = t[:, 32:].copy(); t = t[:, :32].copy()
if int(os.environ.get("HATB200_C3_PROBE", "0")) & 1:
    pr = t[:, 31][t[:, 31] > 0]
    print("in-kernel probe: cycles per MMA over 64 back-to-back issues: median %.1f min %.1f max %.1f" % (np.median(pr) / 64, pr.min() / 64, pr.max() / 64))
t[:, 31] = 0
t0 = t[live][t[live] > 0].min()
print(Cin, N, "act", ACT, "ctas", live.sum(), "span us", (t[live].max() - t0) / 1000)
for cta in list(np.nonzero(live)[0][:2]) + [int(np.nonzero(live)[0][-1])]:
    r_ = t[cta]
    ev = {k: int(r_[k] - t0) for k in range(32) if r_[k] > 0}
    if r_[3] > r_[2] > 0:
        print("   SM clock during the first tile's MMAs: %.3f GHz" % ((clk[cta][3] - clk[cta][2]) / float(r_[3] - r_[2])))
    print("cta", cta, "MMA issue us per tile:", [round((r_[4 * i + 3] - r_[4 * i + 2]) / 1000, 2) for i in range(4) if r_[4 * i + 3] > 0])
    print("cta", cta, {k: round(v / 1000, 2) for k, v in sorted(ev.items())})
