"""Engineering: one ffn2_tc launch at the bench shape (for ncu)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from super_resolution_b200 import _lib  # noqa: E402

lib = _lib.lib()
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
for _ in range(3):
    _lib.check(lib.hatb200_ffn_tail_fused(p(h1), 768, B, H, W, hid, p(dw_w), p(dw_b), p(w2), p(b2), N, p(resid), p(out),
                                          p(gamma), p(beta), p(ln), 192, None))
torch.cuda.synchronize()
print("ok")
