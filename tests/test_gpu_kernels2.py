"""GPU: isolated parity of the kernels that round 1 only covered through whole-forward goldens:
the tcgen05 Toeplitz 13x13 kernel (plk_tc), conv_first, the conv_last tensor-core route, hab_finish, and a BIT-EXACT
integer test of the pixel-shuffle TMA store -- all through the C ABI, against torch-CPU fp32 restatements."""
import ctypes as C

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import hat_oracle as O
from super_resolution_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MEAN = (0.4488, 0.4371, 0.4040, 0.0)


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _tok(x):      # (B,C,H,W) -> (B*H*W, C)
    return x.permute(0, 2, 3, 1).reshape(-1, x.shape[1]).contiguous()


def _img(t, B, H, W):
    return t.view(B, H, W, -1).permute(0, 3, 1, 2)


@pytest.mark.parametrize("B,H,W,with_dyn", [(1, 16, 32, True), (2, 32, 32, True), (3, 64, 64, True), (1, 48, 96, False),
                                            (16, 64, 64, True)])
def test_esc_plk_tcgen05(B, H, W, with_dyn):
    """esc_arch.py:119-123: conv13x13 on the first 16 channels + per-sample dynamic depthwise 3x3."""
    g = torch.Generator().manual_seed(31 + B)
    x = torch.randn(B, 16, H, W, generator=g).bfloat16().float()
    plk = (torch.randn(16, 16, 13, 13, generator=g) * 0.02)
    dyn = torch.randn(B, 16, 9, generator=g) * 0.2 if with_dyn else None
    ref = F.conv2d(x, plk.bfloat16().float(), padding=6)
    if with_dyn:
        ref = ref + torch.cat([F.conv2d(x[b:b + 1], dyn[b].view(16, 1, 3, 3), padding=1, groups=16) for b in range(B)])
    x16 = _tok(x).to(DEV).bfloat16()
    out = torch.full((B * H * W, 64), 3.0, dtype=torch.bfloat16, device=DEV)
    plk_d, dyn_d = plk.to(DEV), (dyn.to(DEV) if with_dyn else None)        # keep the device copies alive over the call
    _lib.check(_lib.lib().hatb200_esc_plk(_p(x16), B, H, W, _p(plk_d), _p(dyn_d), _p(out), 64, _stream()))
    torch.cuda.synchronize()
    got = _img(out[:, :16].float().cpu(), B, H, W)
    err = (got - ref).abs().max().item()
    assert err <= 2.0 ** -7 * max(1.0, ref.abs().max().item()), err
    assert (out[:, 16:] == 3.0).all()                       # only channels [0, 16) are written


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,H,W,Cn,rng", [(2, 16, 32, 180, 1.0), (1, 32, 16, 144, 255.0)])
def test_conv_first(dtype, B, H, W, Cn, rng):
    g = torch.Generator().manual_seed(41)
    x = torch.rand(B, 3, H, W, generator=g)
    w = torch.randn(Cn, 3, 3, 3, generator=g) * 0.2
    b = torch.randn(Cn, generator=g) * 0.1
    xin = x.to(dtype).float()
    mean = torch.tensor(MEAN[:3]).view(1, 3, 1, 1)
    ref = _tok(F.conv2d((xin - mean) * rng, w, b, padding=1))
    out = torch.zeros(B * H * W, Cn, device=DEV)
    m4 = (C.c_float * 4)(*MEAN)
    xd, wd, bd = x.to(DEV).to(dtype), w.to(DEV), b.to(DEV)
    _lib.check(_lib.lib().hatb200_conv_first(_p(xd), _lib.dtype_code(dtype), B, 3, H, W, _p(wd), _p(bd), Cn, m4, rng,
                                             _p(out), _stream()))
    torch.cuda.synchronize()
    err = (out.cpu() - ref).abs().max().item()
    assert err <= 1e-5 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("use_tc,ydt", [(1, torch.bfloat16), (1, torch.float32), (2, torch.bfloat16), (2, torch.float32),
                                        (0, torch.bfloat16), (0, torch.float32)])
@pytest.mark.parametrize("B,H,W,rng", [(1, 32, 64, 1.0), (2, 128, 128, 255.0)])
def test_conv_last_routes(use_tc, ydt, B, H, W, rng):
    """hat_arch.py:856-858, the routes of the bf16 forward: 1 = tensor-core N=3->16 with the affine + NCHW store in the
    conv3 epilogue (one pass), 2 = the same conv followed by the NCHW/affine pass (round 1), 0 = CUDA-core kernel."""
    g = torch.Generator().manual_seed(43)
    x = torch.randn(B, 64, H, W, generator=g).bfloat16().float()
    w = torch.randn(3, 64, 3, 3, generator=g) * 0.05
    b = torch.randn(3, generator=g) * 0.1
    mean = torch.tensor(MEAN[:3]).view(1, 3, 1, 1)
    wr = w.bfloat16().float() if use_tc else w
    ref = F.conv2d(x, wr, b, padding=1) / rng + mean
    y = torch.zeros(B, 3, H, W, dtype=ydt, device=DEV)
    m4 = (C.c_float * 4)(*MEAN)
    xd, wd, bd = _tok(x).to(DEV).bfloat16(), w.to(DEV), b.to(DEV)
    _lib.check(_lib.lib().hatb200_conv_last(_p(xd), _lib.BF16, B, H, W, _p(wd), _p(bd), 3, m4, rng, _p(y),
                                            _lib.dtype_code(ydt), use_tc, _stream()))
    torch.cuda.synchronize()
    err = (y.float().cpu() - ref).abs().max().item()
    tol = 1e-5 if ydt == torch.float32 else 2.0 ** -8
    assert err <= tol * max(1.0, ref.abs().max().item()), err


def test_conv_last_u8_mode_is_tensor2img_of_the_bf16_image():
    """y_dtype U8: (B, H, W, 3) bytes, BGR = tensor2img (img_util.py:38-95) of the bf16 NCHW image of the same entry."""
    from super_resolution_b200 import tensor2img
    g = torch.Generator().manual_seed(44)
    B, H, W = 2, 64, 96
    x = torch.randn(B, 64, H, W, generator=g).bfloat16().float()
    w = torch.randn(3, 64, 3, 3, generator=g) * 0.02
    b = torch.rand(3, generator=g)                     # values spread over and beyond [0, 1]: the clamp matters
    m4 = (C.c_float * 4)(*MEAN)
    xd, wd, bd = _tok(x).to(DEV).bfloat16(), w.to(DEV), b.to(DEV)
    y16 = torch.zeros(B, 3, H, W, dtype=torch.bfloat16, device=DEV)
    y8 = torch.zeros(B, H, W, 3, dtype=torch.uint8, device=DEV)
    for y in (y16, y8):
        _lib.check(_lib.lib().hatb200_conv_last(_p(xd), _lib.BF16, B, H, W, _p(wd), _p(bd), 3, m4, 1.0, _p(y),
                                                _lib.dtype_code(y.dtype), 1, _stream()))
    torch.cuda.synchronize()
    assert y16.float().min() < 0 and y16.float().max() > 1
    want = torch.stack([torch.from_numpy(tensor2img(y16[i:i + 1])) for i in range(B)])
    assert torch.equal(y8.cpu(), want)


@pytest.mark.parametrize("B,Cn,hw", [(1, 180, 1024), (3, 144, 4096), (16, 180, 4096)])
def test_hab_finish(B, Cn, hw):
    """ECA gate (hat_arch.py:66-78) and ESC dynamic kernel (esc_arch.py:95-100) from per-tile partial sums."""
    g = torch.Generator().manual_seed(47)
    nchunk, pdim = hw // 128, 16
    ep = torch.randn(B, nchunk, Cn, generator=g) * 10
    dp = torch.randn(B, nchunk, pdim, generator=g) * 10
    w5 = torch.randn(5, generator=g)
    w1, b1 = torch.randn(8, 16, generator=g) * 0.3, torch.randn(8, generator=g) * 0.1
    w2, b2 = torch.randn(144, 8, generator=g) * 0.3, torch.randn(144, generator=g) * 0.1
    gap = ep.sum(1) / hw                                                # (B, C)
    ref_s = torch.sigmoid(F.conv1d(gap.view(B, 1, Cn), w5.view(1, 1, 5), padding=2)).view(B, Cn)
    ref_d = (F.gelu((dp.sum(1) / hw) @ w1.t() + b1) @ w2.t() + b2).view(B, pdim, 9)
    s = torch.zeros(B, Cn, device=DEV)
    d = torch.zeros(B, pdim, 9, device=DEV)
    dev = [t.to(DEV) for t in (ep, w5, dp, w1, b1, w2, b2)]
    _lib.check(_lib.lib().hatb200_hab_finish(_p(dev[0]), nchunk, B, Cn, hw, _p(dev[1]), _p(s), _p(dev[2]), pdim,
                                             _p(dev[3]), _p(dev[4]), _p(dev[5]), _p(dev[6]), _p(d), _stream()))
    torch.cuda.synchronize()
    assert (s.cpu() - ref_s).abs().max().item() <= 1e-6
    assert (d.cpu() - ref_d).abs().max().item() <= 1e-5 * max(1.0, ref_d.abs().max().item())


@pytest.mark.parametrize("r,B,H,W", [(2, 1, 16, 16), (2, 3, 32, 48), (3, 2, 16, 32), (2, 1, 128, 128), (3, 1, 64, 64)])
def test_pixel_shuffle_store_is_bit_exact(r, B, H, W):
    """The PixelShuffle(r) permutation folded into the 5-D TMA store of conv3_tc (hat_arch.py:593-605) is integer
    address arithmetic: out[b, y r + i, x r + j, c] must hold conv channel c r^2 + i r + j of pixel (b, y, x), exactly.
    Every run makes the conv output a small integer that names one coordinate (exact in bf16)."""
    N = 64 * r * r
    src = O.pixel_shuffle_source(64, H, W, r)                 # (64, H r, W r): flat index into (64 r^2, H, W)
    n_of, rem = src // (H * W), src % (H * W)
    y_of, x_of = rem // W, rem % W
    yy, xx, bb = torch.meshgrid(torch.arange(H), torch.arange(W), torch.arange(B), indexing="ij")

    def run(pix_code, bias):
        a = torch.zeros(B * H * W, 64)
        if pix_code is not None:
            a[:, 0] = pix_code.permute(2, 0, 1).reshape(-1).float()          # (b, y, x) token order
        w = torch.zeros(N, 64, 3, 3)
        if pix_code is not None:
            w[:, 0, 1, 1] = 1.0
        out = torch.full((B, H * r, W * r, 64), -1.0, dtype=torch.bfloat16, device=DEV)
        ad, wd, bd = a.to(DEV).bfloat16(), w.to(DEV), bias.float().to(DEV)
        _lib.check(_lib.lib().hatb200_conv_shuffle(_p(ad), B, H, W, _p(wd), _p(bd), r, _p(out), _stream()))
        torch.cuda.synchronize()
        return out.float().cpu().permute(0, 3, 1, 2).numpy()                # (B, 64, H r, W r)

    zero_b = torch.zeros(N)
    got_x = run(xx, zero_b)
    got_y = run(yy, zero_b)
    got_b = run(bb, zero_b)
    got_nlo = run(None, torch.arange(N) % 256)
    got_nhi = run(None, torch.arange(N) // 256)
    for b in range(B):
        assert np.array_equal(got_x[b], x_of.astype(np.float32)), "x coordinate of the source pixel"
        assert np.array_equal(got_y[b], y_of.astype(np.float32)), "y coordinate of the source pixel"
        assert np.array_equal(got_b[b], np.full_like(got_b[b], b)), "sample of the source pixel"
        assert np.array_equal(got_nlo[b] + 256 * got_nhi[b], n_of.astype(np.float32)), "conv channel c r^2 + i r + j"


def _ffn_tail_case(B, H, W, hid, N, seed):
    g = torch.Generator().manual_seed(seed)
    in_pitch, ln_pitch = ((2 * hid + 63) // 64) * 64, ((N + 63) // 64) * 64
    h1 = torch.randn(B, 2 * hid, H, W, generator=g).bfloat16().float()
    dw_w = torch.randn(2 * hid, 1, 3, 3, generator=g) * 0.3
    dw_b = torch.randn(2 * hid, generator=g) * 0.1
    w2 = torch.randn(N, hid, generator=g) * (1.0 / hid ** 0.5)
    b2 = torch.randn(N, generator=g) * 0.1
    resid = torch.randn(B * H * W, N, generator=g)
    gamma, beta = 1.0 + 0.1 * torch.randn(N, generator=g), 0.02 * torch.randn(N, generator=g)
    t = F.conv2d(h1, dw_w, dw_b, padding=1, groups=2 * hid)
    gated = (t[:, :hid] * F.silu(t[:, hid:])).bfloat16().float()               # G is a bf16 tensor-core operand
    v = _tok(gated) @ w2.bfloat16().float().t() + b2 + resid
    ln = F.layer_norm(v, (N,), gamma, beta, 1e-5)
    h1t = torch.full((B * H * W, in_pitch), float("nan"), dtype=torch.bfloat16, device=DEV)   # pad columns: never read
    h1t[:, :2 * hid] = _tok(h1).to(DEV).bfloat16()
    out = torch.zeros(B * H * W, N, device=DEV)
    ln_out = torch.full((B * H * W, ln_pitch), 5.0, dtype=torch.bfloat16, device=DEV)
    dev = [x.to(DEV) for x in (dw_w, dw_b, w2, b2, resid, gamma, beta)]
    _lib.check(_lib.lib().hatb200_ffn_tail_fused(_p(h1t), in_pitch, B, H, W, hid, _p(dev[0]), _p(dev[1]), _p(dev[2]),
                                                 _p(dev[3]), N, _p(dev[4]), _p(out), _p(dev[5]), _p(dev[6]), _p(ln_out),
                                                 ln_pitch, _stream()))
    torch.cuda.synchronize()
    return out.cpu(), v, ln_out.float().cpu(), ln, N


@pytest.mark.parametrize("B,H,W,hid,N", [(1, 16, 8, 360, 180), (2, 32, 32, 360, 180), (1, 48, 24, 288, 144),
                                         (3, 16, 16, 96, 48), (16, 64, 64, 360, 180), (1, 256, 256, 360, 180)])
def test_ffn_tail_fused_tcgen05(B, H, W, hid, N):
    """hat_arch.py:112-119 + :237: dw3x3 + gate -> fc2 -> + residual -> LayerNorm, one kernel; vs torch fp32 on the same
    bf16-rounded operands (tolerance: the dwgate / linear_fused tests' bf16 bar on the gated operand, fp32 elsewhere)."""
    out, v, ln_out, ln, N = _ffn_tail_case(B, H, W, hid, N, 50 + B)
    err = (out - v).abs().max().item()
    assert err <= 2.0 ** -7 * max(1.0, v.abs().max().item()), f"fp32 stream max-abs {err:.3e}"
    err_ln = (ln_out[:, :N] - ln).abs().max().item()
    assert err_ln <= 2.0 ** -6 * max(1.0, ln.abs().max().item()), f"LayerNorm output max-abs {err_ln:.3e}"
    assert ln_out[:, N:].abs().max().item() == 0 if ln_out.shape[1] > N else True       # K padding of the consumer
    # run to run identical (no race between the dw warps, the MMA thread and the epilogue)
    out2, _, ln2, _, _ = _ffn_tail_case(B, H, W, hid, N, 50 + B)
    assert torch.equal(out, out2) and torch.equal(ln_out, ln2)
