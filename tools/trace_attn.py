# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""Cycle budget of one softmax warp per CTA of the OCAB attention kernel (engineering aid).
Slots: 0 wait Q, 1 wait S, 2 first TMEM load, 3 wait P buffer, 4 emit, 5 later TMEM loads, 6 guard + arrives, 7 flush."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HATB200_GEMM_TRACE"] = "1"
os.environ["HATB200_ATTN_TRACE"] = "1"
import numpy as np, torch
from super_resolution_b200 import _lib
lib = C.CDLL(_lib.LIB_PATH)
L = _lib.lib()
p = lambda t: C.c_void_p(t.data_ptr())
B, H, W, Cn, heads = 16, 64, 64, 180, 6
qkv = torch.randn(B * H * W, 576, device="cuda").bfloat16()
out = torch.zeros(B * H * W, 192, dtype=torch.bfloat16, device="cuda")
table = torch.randn(heads, 39 * 39, device="cuda") * 0.5
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    _lib.check(L.hatb200_ocab_attention(p(qkv), _lib.BF16, 576, B, H, W, Cn, heads, 16, 24, p(table), p(out), 192, 1, st))
torch.cuda.synchronize()
buf = np.zeros((1024, 64), dtype=np.uint64)
lib.hatb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_void_p]
lib.hatb200_debug_gemm_trace(buf.ctypes.data, None)
t = buf.astype(np.int64)
live = t[:, 1] > 0
names = ["waitQ", "waitS", "ld0", "waitPbuf", "emit", "ldN", "arrive", "flush"]
for off, nm in ((0, "warp2"), (16, "warp6")):
    m = t[live, off:off + 8].mean(axis=0)
    print(nm, "total kcycles %.1f" % (m.sum() / 1e3), " ".join(f"{n}={v / 1e3:.1f}" for n, v in zip(names, m)))
