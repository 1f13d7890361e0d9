# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""Per-CTA timeline of the one-pass conv_last launch at the bench shape (16 x 256 x 256 HR pixels, 64 -> 3).
Slots as in tools/trace_conv3.py.   python tools/trace_conv_last.py [use_tc: 1 one pass, 2 two pass]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HATB200_GEMM_TRACE"] = "1"
import numpy as np, torch
from super_resolution_b200 import _lib
lib = C.CDLL(_lib.LIB_PATH)
L = _lib.lib()
p = lambda t: C.c_void_p(t.data_ptr())
mode = int(sys.argv[1]) if len(sys.argv) > 1 else 1
B, H, W = 16, 256, 256
x = torch.randn(B * H * W, 64, device="cuda").bfloat16()
w = torch.randn(3, 64, 3, 3, device="cuda") * 0.05
b = torch.randn(3, device="cuda")
y = torch.zeros(B, 3, H, W, dtype=torch.bfloat16, device="cuda")
m4 = (C.c_float * 4)(0.4488, 0.4371, 0.4040, 0.0)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    _lib.check(L.hatb200_conv_last(p(x), _lib.BF16, B, H, W, p(w), p(b), 3, m4, 1.0, p(y), _lib.BF16, mode, st))
torch.cuda.synchronize()
buf = np.zeros((1024, 64), dtype=np.uint64)
lib.hatb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_void_p]
lib.hatb200_debug_gemm_trace(buf.ctypes.data, None)
t = buf.astype(np.int64)[:, :32]
live = t[:, 0] > 0
t0 = t[live][t[live] > 0].min()
print("mode", mode, "ctas", live.sum(), "first-4-tile span us", (t[live].max() - t0) / 1000)
for cta in (0, 2):
    r_ = t[cta]
    print("cta", cta, {k: round(int(r_[k] - t0) / 1000, 2) for k in range(32) if r_[k] > 0})
