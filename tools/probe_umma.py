# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""Probe tcgen05 A-descriptor addressing with unaligned start rows / arbitrary SBO (see csrc/tc_probe.cu)."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from super_resolution_b200 import _lib

lib = C.CDLL(_lib.LIB_PATH)
lib.hatb200_probe_umma.restype = C.c_int
lib.hatb200_probe_umma.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
R = 256
a = torch.zeros(R, 64)
a[:, 0::2] = torch.arange(R, dtype=torch.float32)[:, None]
a[:, 1::2] = torch.arange(1, 64, 2, dtype=torch.float32)[None, :]
a = a.bfloat16().cuda()
b = torch.eye(64).bfloat16().cuda()
out = torch.zeros(128, 64, device="cuda")
for r0, sbo, bo in [(0, 1024, 0), (1, 1024, 0), (1, 1024, 1), (3, 1024, 0), (3, 1024, 3), (8, 1024, 0), (0, 1280, 0),
                    (0, 1280, 2), (1, 1280, 0), (1, 1280, 1), (0, 1152, 0), (0, 2432, 0), (5, 2432, 0), (5, 2432, 5)]:
    out.zero_()
    rc = lib.hatb200_probe_umma(a.data_ptr(), R, b.data_ptr(), r0, sbo, bo, out.data_ptr(), None)
    torch.cuda.synchronize()
    o = out.cpu()
    rows = o[:, 0].int().tolist()
    exp = [r0 + (i // 8) * (sbo // 128) + (i % 8) for i in range(128)]
    ok_rows = sum(1 for i in range(128) if exp[i] < R and rows[i] == exp[i])
    n_exp = sum(1 for i in range(128) if exp[i] < R)
    # chunk order: odd columns should equal their own index; even columns all equal the row id
    good_cols = 0
    for i in range(128):
        if exp[i] < R:
            good_cols += int(torch.equal(o[i, 1::2], torch.arange(1, 64, 2, dtype=torch.float32)) and
                             bool((o[i, 0::2] == o[i, 0]).all()))
    print(f"r0={r0} sbo={sbo} bo={bo} rc={rc}: rows ok {ok_rows}/{n_exp}, rows with intact chunk order {good_cols}/{n_exp};"
          f" first rows {rows[:10]} r8.. {rows[8:12]}")
