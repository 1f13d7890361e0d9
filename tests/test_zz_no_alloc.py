"""GPU: hatb200_dwconv_gate_noalloc (include/hatb200.h; csrc/noalloc.cu) -- the entry point for callers that
cannot call cudaMalloc -- against the allocating hatb200_dwconv_gate, through the C ABI, on two calls in a
row (the second with another width, so contents left in the shared static scratch by the first would show).

Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

from super_resolution_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _case(dtype, B, H, W, hid, seed):
    in_pitch, out_pitch = ((2 * hid + 63) // 64) * 64, ((hid + 63) // 64) * 64
    g = torch.Generator().manual_seed(seed)
    x = torch.zeros(B * H * W, in_pitch)
    x[:, :2 * hid] = torch.randn(B * H * W, 2 * hid, generator=g)
    w = (torch.randn(2 * hid, 1, 3, 3, generator=g) * 0.3).to(DEV)
    b = (torch.randn(2 * hid, generator=g) * 0.1).to(DEV)
    xt = x.to(DEV).to(dtype)
    outs = []
    for name in ("hatb200_dwconv_gate", "hatb200_dwconv_gate_noalloc"):
        out = torch.full((B * H * W, out_pitch), 3.0, dtype=dtype, device=DEV)
        fn = getattr(_lib.lib(), name)
        _lib.check(fn(_p(xt), _lib.dtype_code(dtype), in_pitch, B, H, W, hid, _p(w), _p(b), _p(out), out_pitch,
                      _stream()))
        torch.cuda.synchronize()
        outs.append(out.float().cpu())
    return outs


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_matches_allocating_entry_twice(dtype):
    # same kernels, same operands: the two entries must agree to the last bit; the tolerance the other
    # dwconv_gate test uses (1e-5 fp32, 2^-7 bf16, relative to max |ref|) is the stated bar
    tol = 1e-5 if dtype == torch.float32 else 2.0 ** -7
    for call, (B, H, W, hid, seed) in enumerate([(2, 16, 32, 360, 5), (1, 16, 16, 96, 6)]):
        ref, got = _case(dtype, B, H, W, hid, seed)
        assert torch.isfinite(got).all()
        err = (got - ref).abs().max().item()
        assert err <= tol * max(1.0, ref.abs().max().item()), f"call {call}: max-abs {err:.3e}"
        assert ref[:, :hid].abs().max().item() > 0.1        # the comparison is not zero against zero


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    hid = _lib.NOALLOC_MAX_HID + 8
    x = torch.zeros(16 * 16, 2 * hid, device=DEV)
    w = torch.zeros(2 * hid, 1, 3, 3, device=DEV)
    b = torch.zeros(2 * hid, device=DEV)
    out = torch.zeros(16 * 16, hid, device=DEV)
    rc = _lib.lib().hatb200_dwconv_gate_noalloc(_p(x), _lib.F32, 2 * hid, 1, 16, 16, hid, _p(w), _p(b), _p(out), hid,
                                                _stream())
    assert rc == -1                                          # HATB200_EINVAL
    msg = _lib.lib().hatb200_last_error().decode()
    assert "static weight scratch" in msg and str(_lib.NOALLOC_MAX_HID) in msg
