"""GPU: single-kernel parity through the C ABI entry points (include/hatb200.h) against torch-CPU fp32
restatements of the same op (the oracle's building blocks)."""
import ctypes as C

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import hat_oracle as O
from super_resolution_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _tokens(x_bchw, pitch, dtype):
    """(B,C,H,W) cpu fp32 -> channels-last tokens (B*H*W, pitch) on device, zero padded."""
    b, c, h, w = x_bchw.shape
    t = torch.zeros(b * h * w, pitch, dtype=dtype, device=DEV)
    t[:, :c] = x_bchw.permute(0, 2, 3, 1).reshape(-1, c).to(DEV).to(dtype)
    return t


def _conv_case(B, H, W, Cin, N, k, dtype, use_tc, act=0, coff=0, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, Cin + coff, H, W, generator=g)
    w = torch.randn(N, Cin, k, k, generator=g) * (1.0 / (Cin * k * k) ** 0.5)
    bias = torch.randn(N, generator=g) * 0.1
    pitch = Cin + coff if Cin + coff == 16 else ((Cin + coff + 63) // 64) * 64
    opitch = ((N + 63) // 64) * 64
    xin = x.bfloat16().float() if dtype == torch.bfloat16 else x
    win = w.bfloat16().float() if (dtype == torch.bfloat16 and use_tc) else w
    ref = F.conv2d(xin[:, coff:], win, bias, padding=k // 2)
    if act == 1:
        ref = F.gelu(ref)
    elif act == 2:
        ref = F.leaky_relu(ref, 0.01)
    xt = _tokens(x, pitch, dtype)
    out = torch.full((B * H * W, opitch), 7.0, dtype=dtype, device=DEV)
    wd, bd = w.to(DEV), bias.to(DEV)
    rc = _lib.lib().hatb200_conv_tokens(_p(xt), _lib.dtype_code(dtype), pitch, coff, Cin, B, H, W, _p(wd), _p(bd), N,
                                        k, k, act, _p(out), opitch, use_tc, _stream())
    _lib.check(rc)
    torch.cuda.synchronize()
    got = out[:, :N].float().cpu().view(B, H, W, N).permute(0, 3, 1, 2)
    return got, ref


CONV_SHAPES = [
    # B, H, W, Cin, N, k   (the shapes of HAT's layers)
    (1, 16, 16, 180, 180, 1),     # esc aggr / proj / q_proj
    (2, 16, 32, 180, 720, 1),     # ffn fc1 (3 n-tiles of 240)
    (1, 16, 16, 360, 180, 1),     # ffn fc2
    (1, 32, 16, 180, 540, 1),     # fused q/kv projection (3 n-tiles of 192)
    (1, 16, 32, 180, 60, 3),      # CAB conv1
    (2, 16, 16, 60, 180, 3),      # CAB conv2
    (1, 32, 32, 180, 180, 3),     # RHAG conv / conv_after_body
    (1, 16, 16, 64, 256, 3),      # upsample conv
    (1, 16, 16, 180, 64, 3),      # conv_before_upsample
    (2, 32, 16, 16, 16, 13),      # ESC large-kernel conv on the 16-channel slice
    (1, 16, 16, 144, 6, 3),       # HAT-S CAB conv1 (mid = 6)
    (1, 16, 16, 6, 144, 3),       # HAT-S CAB conv2
    (1, 16, 16, 48, 96, 1),       # small dims
]


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_tokens_fp32_cuda_core(shape):
    got, ref = _conv_case(*shape, torch.float32, 0, act=1)
    assert (got - ref).abs().max().item() <= 2e-5 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("shape", CONV_SHAPES)
@pytest.mark.parametrize("use_tc", [0, 1])
def test_conv_tokens_bf16(shape, use_tc):
    got, ref = _conv_case(*shape, torch.bfloat16, use_tc, act=2)
    # inputs are bf16-exact in the reference; the only error is fp32 accumulation order + bf16 output rounding
    tol = 2.0 ** -8 * max(1.0, ref.abs().max().item())
    assert (got - ref).abs().max().item() <= tol


def test_conv_tokens_channel_offset_tc():
    got, ref = _conv_case(1, 16, 16, 64, 64, 3, torch.bfloat16, 1, coff=64)
    assert (got - ref).abs().max().item() <= 2.0 ** -8 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_layernorm(dtype):
    M, Cn = 512, 180
    g = torch.Generator().manual_seed(1)
    x = torch.randn(M, Cn, generator=g) * 3 + 1.5
    gamma, beta = torch.randn(Cn, generator=g), torch.randn(Cn, generator=g)
    ref = F.layer_norm(x, (Cn,), gamma, beta, 1e-5)
    out = torch.full((M, 192), 7.0, dtype=dtype, device=DEV)
    head = torch.zeros(M, 16, dtype=dtype, device=DEV)
    xd, gd, bd = x.to(DEV), gamma.to(DEV), beta.to(DEV)
    _lib.check(_lib.lib().hatb200_layernorm(_p(xd), M, Cn, _p(gd), _p(bd), _p(out), _lib.dtype_code(dtype), 192,
                                            _p(head), 16, _stream()))
    torch.cuda.synchronize()
    tol = 1e-5 if dtype == torch.float32 else 2.0 ** -7
    assert (out[:, :Cn].float().cpu() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
    assert out[:, Cn:].abs().max().item() == 0              # K padding of the consumer is zero
    assert torch.equal(head.cpu(), out[:, :16].cpu())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_dwconv_gate(dtype):
    B, H, W, hid = 2, 16, 32, 360
    g = torch.Generator().manual_seed(2)
    x = torch.randn(B, 2 * hid, H, W, generator=g)
    w = torch.randn(2 * hid, 1, 3, 3, generator=g) * 0.3
    b = torch.randn(2 * hid, generator=g) * 0.1
    xin = x.bfloat16().float() if dtype == torch.bfloat16 else x
    t = F.conv2d(xin, w, b, padding=1, groups=2 * hid)
    ref = t[:, :hid] * F.silu(t[:, hid:])
    xt = _tokens(x, 768, dtype)
    out = torch.zeros(B * H * W, 384, dtype=dtype, device=DEV)
    wd, bd = w.to(DEV), b.to(DEV)
    _lib.check(_lib.lib().hatb200_dwconv_gate(_p(xt), _lib.dtype_code(dtype), 768, B, H, W, hid, _p(wd), _p(bd),
                                              _p(out), 384, _stream()))
    torch.cuda.synchronize()
    got = out[:, :hid].float().cpu().view(B, H, W, hid).permute(0, 3, 1, 2)
    tol = 1e-5 if dtype == torch.float32 else 2.0 ** -7
    assert (got - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


def _attention_reference(qkv, B, H, W, Cn, heads, ws, ows, table_compact):
    """OCAB core (hat_arch.py:354-388) on (B,H,W,3C) with pre-scaled q, unfold zero padding, unmasked keys."""
    d = Cn // heads
    q, k, v = qkv[..., :Cn], qkv[..., Cn:2 * Cn], qkv[..., 2 * Cn:]
    nwy, nwx = H // ws, W // ws
    qw = q.view(B, nwy, ws, nwx, ws, Cn).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, Cn)
    kv = torch.cat([k, v], -1).permute(0, 3, 1, 2)
    u = F.unfold(kv, kernel_size=ows, stride=ws, padding=(ows - ws) // 2)
    u = u.view(B, 2, Cn, ows * ows, nwy * nwx).permute(1, 0, 4, 3, 2).reshape(2, -1, ows * ows, Cn)
    nb = qw.shape[0]
    qh = qw.view(nb, -1, heads, d).permute(0, 2, 1, 3)
    kh = u[0].view(nb, -1, heads, d).permute(0, 2, 1, 3)
    vh = u[1].view(nb, -1, heads, d).permute(0, 2, 1, 3)
    T = ws + ows - 1
    qi, ki = torch.arange(ws * ws), torch.arange(ows * ows)
    slot = ((ki // ows)[None, :] - (qi // ws)[:, None] + ws - 1) * T + ((ki % ows)[None, :] - (qi % ws)[:, None] + ws - 1)
    bias = table_compact[:, slot]                                    # (heads, nq, nk)
    a = torch.softmax(qh @ kh.transpose(-2, -1) + bias[None], -1)
    o = (a @ vh).transpose(1, 2).reshape(nb, ws * ws, Cn)
    return o.view(B, nwy, nwx, ws, ws, Cn).permute(0, 1, 3, 2, 4, 5).reshape(B, H, W, Cn)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("geom", [(1, 16, 16, 180, 6), (2, 32, 48, 180, 6), (1, 32, 32, 144, 6), (1, 16, 32, 48, 6)])
def test_ocab_attention_cuda_core(dtype, geom):
    _attention_case(dtype, geom, 0)


@pytest.mark.parametrize("geom", [(1, 16, 16, 180, 6), (2, 32, 48, 180, 6), (1, 32, 32, 144, 6), (1, 16, 32, 48, 6),
                                  (3, 64, 64, 180, 6)])
def test_ocab_attention_tcgen05(geom):
    """tcgen05 flash-style kernel (tc_attn.cu): bf16 operands, fp32 softmax; P is rounded to bf16 for the PV MMA."""
    _attention_case(torch.bfloat16, geom, 1)


def test_ocab_attention_tcgen05_wide_logit_spread():
    """Logits spread over hundreds of octaves inside one 96-key chunk: the one-pass softmax (reference exponent from
    the chunk's first 32 keys) overflows and must take its exact two-pass path; the result is a near-hard max."""
    _attention_case(torch.bfloat16, (2, 32, 48, 180, 6), 1, qscale=40.0)


def _attention_case(dtype, geom, use_tc, qscale=1.0):
    B, H, W, Cn, heads = geom
    ws, ows = 16, 24
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn(B, H, W, 3 * Cn, generator=g)
    qkv[..., :Cn] *= (Cn // heads) ** -0.5 * qscale
    table = torch.randn(heads, 39 * 39, generator=g) * 0.5
    src = qkv.bfloat16().float() if dtype == torch.bfloat16 else qkv
    ref = _attention_reference(src, B, H, W, Cn, heads, ws, ows, table)
    pitch = ((3 * Cn + 63) // 64) * 64
    t = torch.zeros(B * H * W, pitch, dtype=dtype, device=DEV)
    t[:, :3 * Cn] = qkv.view(-1, 3 * Cn).to(DEV).to(dtype)
    opitch = ((Cn + 63) // 64) * 64
    out = torch.zeros(B * H * W, opitch, dtype=dtype, device=DEV)
    td = table.to(DEV)
    _lib.check(_lib.lib().hatb200_ocab_attention(_p(t), _lib.dtype_code(dtype), pitch, B, H, W, Cn, heads, ws, ows,
                                                 _p(td), _p(out), opitch, use_tc, _stream()))
    torch.cuda.synchronize()
    got = out[:, :Cn].float().cpu().view(B, H, W, Cn)
    tol = 2e-5 if dtype == torch.float32 else 2.0 ** -7
    assert (got - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


def test_tile_helpers_bit_exact():
    B, Cn, h, w = 2, 3, 37, 45
    x = torch.rand(B, Cn, h, w, device=DEV)
    ph, pw = O.mod_pad(h, w, 16)
    dst = torch.empty(B, Cn, h + ph, w + pw, device=DEV)
    _lib.check(_lib.lib().hatb200_reflect_pad(_p(x), _lib.F32, B, Cn, h, w, _p(dst), ph, pw, _stream()))
    assert torch.equal(dst, F.pad(x, (0, pw, 0, ph), "reflect"))
    big = torch.zeros(B, Cn, 64, 80, device=DEV)
    _lib.check(_lib.lib().hatb200_crop_place(_p(dst), _lib.F32, B, Cn, h + ph, w + pw, 5, 7, _p(big), 64, 80, 11, 13,
                                             20, 30, _stream()))
    want = torch.zeros_like(big)
    want[:, :, 11:31, 13:43] = dst[:, :, 5:25, 7:37]
    assert torch.equal(big, want)


# ---- persistent tcgen05 GEMM with the fused epilogue (csrc/tc_gemm.cu) ------------------------------------
def _linear_fused_case(B, H, W, Cin, N, resid, eca, ln, head, out_t, act=0, seed=0):
    g = torch.Generator().manual_seed(seed)
    M = B * H * W
    pitch = -(-Cin // 64) * 64
    a = torch.zeros(M, pitch, dtype=torch.bfloat16)
    a[:, :Cin] = torch.randn(M, Cin, generator=g).bfloat16()
    w = (torch.randn(N, Cin, generator=g) * 0.08)
    bias = torch.randn(N, generator=g) * 0.3
    npitch = -(-N // 64) * 64
    r = torch.randn(M, N, generator=g) if resid else None
    u = torch.full((M, npitch), float("nan"), dtype=torch.bfloat16)       # pad columns must never be read as values
    u[:, :N] = torch.randn(M, N, generator=g).bfloat16()
    sc = torch.rand(B, N, generator=g)
    gamma = 1.0 + 0.2 * torch.randn(N, generator=g)
    beta = 0.1 * torch.randn(N, generator=g)
    # reference (fp32 math on the bf16-rounded operands)
    v = a[:, :Cin].float() @ w.bfloat16().float().t() + bias
    if act == 1:
        v = F.gelu(v)
    if eca:
        v = v + 0.01 * sc.repeat_interleave(H * W, 0) * u[:, :N].float()
    if resid:
        v = v + r
    ln_ref = F.layer_norm(v, (N,), gamma, beta, 1e-5)
    d = lambda t: None if t is None else t.to(DEV)
    a_d, w_d, b_d, r_d, u_d, sc_d, g_d, be_d = map(d, (a, w, bias, r, u, sc, gamma, beta))
    out_f32 = torch.full((M, N), float("nan"), device=DEV) if resid else None
    ot = torch.full((M, npitch), float("nan"), dtype=torch.bfloat16, device=DEV) if out_t else None
    lo = torch.full((M, npitch), float("nan"), dtype=torch.bfloat16, device=DEV) if ln else None
    hd = torch.full((M, 16), float("nan"), dtype=torch.bfloat16, device=DEV) if head else None
    cs = torch.full((M // 128, 16), float("nan"), device=DEV) if head else None
    p = lambda t: None if t is None else _p(t)
    n16 = -(-N // 16) * 16
    nt = -(-n16 // 256)
    npad = nt * (-(-(-(-n16 // nt)) // 16) * 16)                           # kernels.cuh tc_npad()
    zero_to = npitch if (out_t and npitch <= npad) else 0
    _lib.check(_lib.lib().hatb200_linear_fused(
        p(a_d), pitch, Cin, B, H, W, p(w_d), p(b_d), N, act, p(r_d), p(u_d) if eca else None, npitch,
        p(sc_d) if eca else None, 0.01, p(out_f32), p(ot), npitch, zero_to, p(g_d), p(be_d), p(lo), npitch,
        p(hd), p(cs), _stream()))
    torch.cuda.synchronize()
    scale = max(1.0, v.abs().max().item())
    if out_f32 is not None:
        assert (out_f32.cpu() - v).abs().max().item() <= 2e-4 * scale
    if ot is not None:
        got = ot.float().cpu()
        assert (got[:, :N] - v).abs().max().item() <= 2.0 ** -7 * scale
        if zero_to:
            assert torch.equal(got[:, N:zero_to], torch.zeros(M, zero_to - N))   # K padding of the consumer
    if lo is not None:
        got = lo.float().cpu()
        assert (got[:, :N] - ln_ref).abs().max().item() <= 2.0 ** -6 * max(1.0, ln_ref.abs().max().item())
        assert torch.equal(got[:, N:], torch.zeros(M, npitch - N))
        if hd is not None:
            assert torch.equal(hd.cpu(), lo[:, :16].cpu())
            want = lo[:, :16].float().view(M // 128, 128, 16).sum(1).cpu()
            assert (cs.cpu() - want).abs().max().item() <= 1e-3 * max(1.0, want.abs().max().item())


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [
    # B, H, W, Cin, N, resid, eca, ln, head, out_t, act
    (2, 16, 16, 180, 720, False, False, False, False, True, 0),     # ffn fc1: 3 n-tiles of 240, output only
    (2, 16, 16, 192, 576, False, False, False, False, True, 0),     # head-padded qkv: 32-column chunks
    (2, 32, 16, 180, 360, False, False, False, False, True, 1),     # OCAB mlp.0 + GELU, zero-padded to 384
    (3, 16, 16, 180, 180, True, True, True, False, False, 0),       # ESC aggr + HAB combine + norm2
    (2, 16, 32, 360, 180, True, False, True, True, False, 0),       # ffn fc2 + residual + next norm1 + head
    (1, 16, 16, 360, 180, True, False, False, False, True, 0),      # OCAB mlp.2 + residual + bf16 copy
    (2, 16, 16, 48, 48, True, True, True, True, False, 0),          # mini model width
    (5, 32, 32, 144, 144, True, False, True, True, False, 0),       # HAT-S width, many tiles per CTA
    # plain bf16 layers with BN % 128 == 0 run four epilogue warpgroups (two per accumulator stage, half the columns each)
    (2, 16, 16, 96, 128, False, False, False, False, True, 0),      # BN 128: halves of 64 columns
    (4, 32, 32, 180, 250, False, False, False, False, True, 1),     # BN 256, ragged N: pad columns of the upper half are zeros
    (3, 16, 32, 64, 512, False, False, False, False, True, 0),      # two n-tiles of 256
])
def test_linear_fused_tcgen05(shape):
    _linear_fused_case(*shape)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [
    # BASELINE configs[1] sizes: 512 token tiles over 148 persistent CTAs (3-4 tiles each, operand ring wraps ~12x per
    # CTA).  The small shapes above give every CTA at most one tile and never caught a ring stage released before its
    # shared-memory loads had landed (0.3 % of the rows took a later chunk's residual at this size).
    (16, 64, 64, 360, 180, True, False, True, True, False, 0),      # ffn fc2 at bench size
    (16, 64, 64, 180, 180, True, True, True, False, False, 0),      # ESC aggr + HAB combine at bench size
    (1, 256, 256, 360, 180, True, False, True, True, False, 0),     # one 256x256 sample (configs[2])
    (16, 64, 64, 180, 720, False, False, False, False, True, 0),    # ffn fc1 at bench size
])
def test_linear_fused_tcgen05_full_size(shape):
    _linear_fused_case(*shape, seed=3)
    _linear_fused_case(*shape, seed=4)      # run to run: the failure was timing dependent


# ---- persistent halo-resident tcgen05 3x3 conv with the fused epilogue (csrc/tc_conv3.cu) ---------------------
def _conv3_fused_case(B, H, W, Cin, N, resid, ln, head, out_t, colsum, act=0, seed=0):
    g = torch.Generator().manual_seed(seed)
    M = B * H * W
    pitch = -(-Cin // 64) * 64
    x = torch.randn(B, Cin, H, W, generator=g)
    a = torch.full((M, pitch), float("nan"), dtype=torch.bfloat16)
    a[:, Cin:] = 0                                       # K padding of the producer; real channels below
    a[:, :Cin] = x.permute(0, 2, 3, 1).reshape(M, Cin).bfloat16()
    w = torch.randn(N, Cin, 3, 3, generator=g) * 0.05
    bias = torch.randn(N, generator=g) * 0.3
    npitch = -(-N // 64) * 64
    r = torch.randn(M, N, generator=g) if resid else None
    gamma = 1.0 + 0.2 * torch.randn(N, generator=g)
    beta = 0.1 * torch.randn(N, generator=g)
    xin = a[:, :Cin].float().view(B, H, W, Cin).permute(0, 3, 1, 2)
    v = F.conv2d(xin, w.bfloat16().float(), bias, padding=1).permute(0, 2, 3, 1).reshape(M, N)
    if act == 1:
        v = F.gelu(v)
    elif act == 2:
        v = F.leaky_relu(v, 0.01)
    if resid:
        v = v + r
    ln_ref = F.layer_norm(v, (N,), gamma, beta, 1e-5)
    d = lambda t: None if t is None else t.to(DEV)
    a_d, w_d, b_d, r_d, g_d, be_d = map(d, (a, w, bias, r, gamma, beta))
    out_f32 = torch.full((M, N), float("nan"), device=DEV) if resid else None
    ot = torch.full((M, npitch), float("nan"), dtype=torch.bfloat16, device=DEV) if out_t else None
    lo = torch.full((M, npitch), float("nan"), dtype=torch.bfloat16, device=DEV) if ln else None
    hd = torch.full((M, 16), float("nan"), dtype=torch.bfloat16, device=DEV) if head else None
    hcs = torch.full((M // 128, 16), float("nan"), device=DEV) if head else None
    cs = torch.full((M // 128, N), float("nan"), device=DEV) if colsum else None
    p = lambda t: None if t is None else _p(t)
    n16 = -(-N // 16) * 16
    nt = -(-n16 // 256)
    per = -(-n16 // nt)
    npad = nt * (-(-per // (64 if n16 > 32 else 16)) * (64 if n16 > 32 else 16))      # kernels.cuh tc_npad()
    zero_to = npitch if (out_t and npitch <= npad) else 0
    _lib.check(_lib.lib().hatb200_conv_fused(
        p(a_d), pitch, Cin, B, H, W, p(w_d), p(b_d), N, 3, act, p(r_d), p(out_f32), p(ot), npitch, zero_to, p(g_d),
        p(be_d), p(lo), npitch, p(hd), p(hcs), p(cs), _stream()))
    torch.cuda.synchronize()
    scale = max(1.0, v.abs().max().item())
    if out_f32 is not None:
        assert (out_f32.cpu() - v).abs().max().item() <= 3e-4 * scale
    if ot is not None:
        got = ot.float().cpu()
        assert (got[:, :N] - v).abs().max().item() <= 2.0 ** -7 * scale
        if zero_to:
            assert torch.equal(got[:, N:zero_to], torch.zeros(M, zero_to - N))
    if lo is not None:
        got = lo.float().cpu()
        assert (got[:, :N] - ln_ref).abs().max().item() <= 2.0 ** -6 * max(1.0, ln_ref.abs().max().item())
        assert torch.equal(got[:, N:], torch.zeros(M, npitch - N))
        if hd is not None:
            # head rows are tokens; column sums are per 8 x 16 pixel tile, tiles numbered row-major inside a sample
            assert torch.equal(hd.cpu(), lo[:, :16].cpu())
            t = lo[:, :16].float().cpu().view(B, H // 16, 16, W // 8, 8, 16).sum((2, 4)).reshape(-1, 16)
            assert (hcs.cpu() - t).abs().max().item() <= 1e-3 * max(1.0, t.abs().max().item())
    if cs is not None:
        t = v.view(B, H // 16, 16, W // 8, 8, N).sum((2, 4)).reshape(-1, N)
        assert (cs.cpu() - t).abs().max().item() <= 1e-3 * max(1.0, t.abs().max().item())


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [
    # B, H, W, Cin, N, resid, ln, head, out_t, colsum, act
    (2, 16, 32, 180, 60, False, False, False, True, False, 1),      # CAB conv1 + GELU
    (2, 32, 16, 60, 180, False, False, False, True, True, 0),       # CAB conv2 + pool partials
    (1, 32, 32, 180, 180, True, True, True, False, False, 0),       # RHAG conv + group residual + next norm1 + head
    (3, 16, 16, 180, 180, True, False, False, True, False, 0),      # conv_after_body + long skip
    (1, 16, 16, 180, 64, False, False, False, True, False, 2),      # conv_before_upsample + LeakyReLU
    (5, 32, 32, 144, 144, True, True, True, False, False, 0),       # HAT-S width, several tiles per CTA
    (1, 64, 64, 64, 256, False, False, False, True, False, 0),      # upsample-sized n-tile (256 columns)
    (2, 32, 32, 128, 96, False, False, False, True, False, 0),      # 2 k-blocks: the looped (non-unrolled) issuer / producer
    (2, 32, 16, 256, 64, False, False, False, True, False, 1),      # 4 k-blocks, GELU
])
def test_conv3_fused_tcgen05(shape):
    _conv3_fused_case(*shape)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [
    # bench-size launches: 512 pixel tiles over 74 CTA pairs (cta_group::2), 3-4 pair iterations each
    (16, 64, 64, 180, 180, True, True, True, False, False, 0),      # RHAG conv + residual ring + LayerNorm + head
    (16, 64, 64, 180, 60, False, False, False, True, False, 1),     # CAB conv1 + GELU
    (16, 64, 64, 60, 180, False, False, False, True, True, 0),      # CAB conv2 + pool partials
    (3, 48, 80, 180, 180, True, True, True, False, False, 0),       # 3 * 3 * 10 = 90 tiles: ragged pair waves
    (1, 16, 24, 180, 60, False, False, False, True, False, 1),      # 3 tiles: odd count -> single-CTA kernel
])
def test_conv3_fused_tcgen05_full_size(shape):
    _conv3_fused_case(*shape, seed=5)


@pytest.mark.gpu
def test_linear_fused_n_split_pair_opt_in(monkeypatch):
    """HATB200_G3_SPLIT=1: the residual + LayerNorm layers on N-split CTA pairs (each CTA half of the row, LayerNorm
    statistics merged across the cluster through distributed shared memory).  Opt-in because it measured slower; the
    numerics are the same epilogue, so the full-size cases must still pass."""
    monkeypatch.setenv("HATB200_G3_SPLIT", "1")
    _linear_fused_case(16, 64, 64, 360, 180, True, False, True, True, False, 0, seed=3)     # ffn fc2 at bench size
    _linear_fused_case(16, 64, 64, 180, 180, True, True, True, False, False, 0, seed=3)     # ESC aggr at bench size
    _linear_fused_case(2, 16, 32, 360, 180, True, False, True, True, False, 0)              # fewer tiles than pairs
