# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""Engineering: per-CTA %globaltimer trace of one ffn2_tc launch at the bench shape (B=16, 64x64, hid 360, C 180)."""
import ctypes as C
import os
import sys

os.environ["HATB200_GEMM_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from super_resolution_b200 import _lib  # noqa: E402

lib = _lib.lib()
lib.hatb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_void_p]
B, H, W, hid, N = 16, 64, 64, 360, 180
M = B * H * W
dev = "cuda:0"
h1 = torch.randn(M, 768, device=dev).bfloat16()
dw_w = torch.randn(2 * hid, 1, 3, 3, device=dev) * 0.3
dw_b = torch.randn(2 * hid, device=dev) * 0.1
w2 = torch.randn(N, hid, device=dev) * 0.05
b2 = torch.randn(N, device=dev) * 0.1
resid = torch.randn(M, N, device=dev)
out = torch.zeros(M, N, device=dev)
gamma, beta = torch.ones(N, device=dev), torch.zeros(N, device=dev)
ln = torch.zeros(M, 192, device=dev, dtype=torch.bfloat16)
p = lambda t: C.c_void_p(t.data_ptr())
for _ in range(2):
    _lib.check(lib.hatb200_ffn_tail_fused(p(h1), 768, B, H, W, hid, p(dw_w), p(dw_b), p(w2), p(b2), N, p(resid), p(out),
                                          p(gamma), p(beta), p(ln), 192, None))
torch.cuda.synchronize()
buf = np.zeros(1024 * 64, dtype=np.uint64)
lib.hatb200_debug_gemm_trace(buf.ctypes.data, None)
tr = buf.reshape(1024, 64)[:148].astype(np.int64)
t0 = tr[:, [0]].min()
for cta in (0, 1, 70, 147):
    r = tr[cta]
    rel = lambda v: (v - t0) / 1e3 if v else float("nan")
    print(f"CTA {cta}: dw tiles " + " ".join(f"[{rel(r[4*i]):.1f}-{rel(r[4*i+1]):.1f}]" for i in range(4)))
    print(f"        epi      " + " ".join(f"[wait {rel(r[16+4*i]):.1f} go {rel(r[17+4*i]):.1f} end {rel(r[18+4*i]):.1f}]" for i in range(4)))
    print(f"        dw wait cycles: staging {r[32]}  g-tile {r[33]}")
end = tr[:, 16:32].max()
print(f"launch span {(end - t0) / 1e3:.1f} us")
