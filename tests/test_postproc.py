"""Output post-processing right after the path and checkpoint files (8(f3)).

CPU: the numpy restatement (oracle/img_oracle.py) against fixtures produced by the unmodified reference
(`oracle/make_golden.py::gen_postproc`, basicsr/utils/img_util.py:38-95, basicsr/metrics/psnr_ssim.py:12-50), and
against the live reference when it is mounted; checkpoint files in the reference's `{params, params_ema}` layout.
GPU: `hatb200_tensor2img` bit-exact against the fixtures and the oracle; `hatb200_psnr_u8` to 1e-9 relative (the sum of
squares is accumulated in a different order than numpy's pairwise sum)."""
import json
import math
import os

import numpy as np
import pytest
import torch

from oracle import img_oracle as IO
from oracle import load_ref
from oracle.make_golden import postproc_inputs
from tests.helpers import GOLDEN


def _gold():
    z = np.load(os.path.join(GOLDEN, "postproc.npz"), allow_pickle=False)
    return z, json.loads(str(z["psnr"]))


def test_oracle_tensor2img_matches_reference_golden():
    z, _ = _gold()
    a, b, gray, wide = postproc_inputs()
    assert np.array_equal(IO.tensor2img(a.numpy()), z["a_bgr"])
    assert np.array_equal(IO.tensor2img(b.numpy()), z["b_bgr"])
    assert np.array_equal(IO.tensor2img(a.numpy(), rgb2bgr=False), z["a_rgb"])
    assert np.array_equal(IO.tensor2img(gray.numpy()), z["gray"])
    assert np.array_equal(IO.tensor2img(wide.numpy(), min_max=(0, 255)), z["wide"])
    # the fixture exercises what it claims: clamped pixels at both ends and .5 ties rounded to even
    assert z["a_bgr"].min() == 0 and z["a_bgr"].max() == 255
    assert z["a_bgr"][0, :8, 0].tolist() == [0, 2, 2, 4, 126, 128, 254, 254]


def test_oracle_psnr_matches_reference_golden():
    z, psnr = _gold()
    for cb in (0, 4):
        for ty in (False, True):
            got = IO.calculate_psnr(z["a_bgr"], z["b_bgr"], cb, ty)
            assert got == pytest.approx(psnr[f"cb{cb}_y{int(ty)}"], rel=1e-12)
    _, _, gray, _ = postproc_inputs()
    got = IO.calculate_psnr(z["gray"], IO.tensor2img((gray * 0.97).numpy()), 2, True)
    assert got == pytest.approx(psnr["gray_cb2_y1"], rel=1e-12)
    assert IO.calculate_psnr(z["a_bgr"], z["a_bgr"], 0) == float("inf")


def test_oracle_ssim_matches_reference_golden():
    z, _ = _gold()
    ssim = json.loads(str(z["ssim"]))
    for cb in (0, 4):
        for ty in (False, True):
            got = IO.calculate_ssim(z["a_bgr"], z["b_bgr"], cb, ty)
            assert got == pytest.approx(ssim[f"cb{cb}_y{int(ty)}"], rel=1e-12), (cb, ty)
    _, _, gray, _ = postproc_inputs()
    g2 = IO.tensor2img((gray * 0.97).numpy())
    assert IO.calculate_ssim(z["gray"], g2, 2, True) == pytest.approx(ssim["gray_cb2_y1"], rel=1e-12)
    assert IO.calculate_ssim(z["gray"], g2, 0, False) == pytest.approx(ssim["gray_cb0_y0"], rel=1e-12)
    assert IO.calculate_ssim(z["a_bgr"], z["a_bgr"], 0) == pytest.approx(1.0, abs=1e-15)


@pytest.mark.skipif(not load_ref.reference_available(), reason="reference not mounted")
def test_oracle_postproc_matches_live_reference():
    load_ref.load_reference_hat()
    from basicsr.metrics.psnr_ssim import calculate_psnr
    from basicsr.utils.img_util import tensor2img
    g = torch.Generator().manual_seed(11)
    x = torch.rand(1, 3, 33, 47, generator=g) * 1.4 - 0.2
    y = x + 0.05 * torch.randn(x.shape, generator=g)
    ix, iy = tensor2img(x), tensor2img(y)
    assert np.array_equal(IO.tensor2img(x.numpy()), ix)
    for ty in (False, True):
        assert IO.calculate_psnr(ix, iy, 3, ty) == pytest.approx(calculate_psnr(ix, iy, 3, test_y_channel=ty), rel=1e-12)
    from basicsr.metrics.psnr_ssim import calculate_ssim
    for ty in (False, True):
        assert IO.calculate_ssim(ix, iy, 3, ty) == pytest.approx(calculate_ssim(ix, iy, 3, test_y_channel=ty), rel=1e-12)


def test_checkpoint_roundtrip_reference_layout(tmp_path):
    """`save_network` writes `{param_key: state_dict}`; `load_network` reads `params_ema` with a fallback to `params`
    and strips `module.` (basicsr/models/base_model.py:207-252, :289-315)."""
    from oracle import hat_oracle as O
    from super_resolution_b200 import HAT, load_network, save_network
    cfg = dict(upscale=2, img_size=32, window_size=16, compress_ratio=3, squeeze_factor=30, overlap_ratio=0.5,
               depths=[1], embed_dim=24, num_heads=[6], mlp_ratio=2, upsampler='pixelshuffle', resi_connection='1conv')
    src = HAT(**cfg)
    src.load_state_dict(O.randomized_state_dict(src.state_dict(), 3))
    path = str(tmp_path / "net_g_latest.pth")
    save_network(src, path)                                   # {'params': ...}
    blob = torch.load(path, weights_only=True)
    assert list(blob.keys()) == ["params"] and list(blob["params"].keys()) == list(src.state_dict().keys())
    dst = HAT(**cfg)
    load_network(dst, path, strict=True, param_key="params_ema")     # falls back to 'params'
    for k, v in src.state_dict().items():
        assert torch.equal(v, dst.state_dict()[k]), k
    # a checkpoint saved from a DataParallel wrapper: 'module.' prefixes, both keys
    wrapped = {"params": {"module." + k: v for k, v in src.state_dict().items()},
               "params_ema": {"module." + k: v * 0 + 1 if v.is_floating_point() else v for k, v in src.state_dict().items()}}
    torch.save(wrapped, path)
    load_network(dst, path, strict=True, param_key="params_ema")
    assert all(bool((v == 1).all()) for v in dst.state_dict().values() if v.is_floating_point())
    # strict=False skips keys whose size differs instead of raising
    bad = {"params": dict(src.state_dict())}
    bad["params"]["conv_last.bias"] = torch.zeros(7)
    torch.save(bad, path)
    with pytest.raises(RuntimeError):
        load_network(HAT(**cfg), path, strict=True)
    load_network(HAT(**cfg), path, strict=False)


def test_postproc_requires_cuda():
    from super_resolution_b200 import calculate_psnr, tensor2img
    with pytest.raises(RuntimeError):
        tensor2img(torch.rand(3, 8, 8))
    with pytest.raises(RuntimeError):
        calculate_psnr(torch.zeros(8, 8, 3, dtype=torch.uint8), torch.zeros(8, 8, 3, dtype=torch.uint8), 0)
    with pytest.raises(TypeError):
        tensor2img("not a tensor")


@pytest.mark.gpu
def test_gpu_tensor2img_bit_exact():
    from super_resolution_b200 import tensor2img
    z, _ = _gold()
    a, b, gray, wide = postproc_inputs()
    dev = "cuda:0"
    assert np.array_equal(tensor2img([a.to(dev)])[0], z["a_bgr"])
    assert np.array_equal(tensor2img(b.to(dev)), z["b_bgr"])
    assert np.array_equal(tensor2img(a.to(dev), rgb2bgr=False), z["a_rgb"])
    assert np.array_equal(tensor2img(gray.to(dev)), z["gray"])
    assert np.array_equal(tensor2img(wide.to(dev), min_max=(0, 255)), z["wide"])
    # HR-sized image against the oracle (ragged size, bf16 input is widened exactly)
    g = torch.Generator().manual_seed(5)
    big = torch.rand(1, 3, 1021, 773, generator=g) * 1.2 - 0.1
    assert np.array_equal(tensor2img(big.to(dev)), IO.tensor2img(big.numpy()))
    b16 = big.bfloat16()
    assert np.array_equal(tensor2img(b16.to(dev)), IO.tensor2img(b16.float().numpy()))
    t = tensor2img(big.to(dev), to_numpy=False)
    assert t.is_cuda and t.dtype == torch.uint8 and tuple(t.shape) == (1021, 773, 3)


@pytest.mark.gpu
def test_gpu_psnr_matches_reference():
    from super_resolution_b200 import calculate_psnr
    z, psnr = _gold()
    dev = "cuda:0"
    ia, ib = torch.from_numpy(z["a_bgr"]).to(dev), torch.from_numpy(z["b_bgr"]).to(dev)
    for cb in (0, 4):
        for ty in (False, True):
            assert calculate_psnr(ia, ib, cb, test_y_channel=ty) == pytest.approx(psnr[f"cb{cb}_y{int(ty)}"], rel=1e-9)
    assert calculate_psnr(ia, ia, 0) == float("inf")
    assert calculate_psnr(ia.permute(2, 0, 1), ib.permute(2, 0, 1), 4, input_order="CHW") == \
        pytest.approx(psnr["cb4_y0"], rel=1e-9)
    _, _, gray, _ = postproc_inputs()
    g2 = torch.from_numpy(IO.tensor2img((gray * 0.97).numpy())).to(dev)
    assert calculate_psnr(torch.from_numpy(z["gray"]).to(dev), g2, 2, test_y_channel=True) == \
        pytest.approx(psnr["gray_cb2_y1"], rel=1e-9)
    # larger than one reduction wave, against the oracle
    g = torch.Generator().manual_seed(9)
    x = torch.randint(0, 256, (1500, 1100, 3), generator=g, dtype=torch.uint8)
    y = (x.int() + torch.randint(-6, 7, x.shape, generator=g)).clamp(0, 255).to(torch.uint8)
    for ty in (False, True):
        ref = IO.calculate_psnr(x.numpy(), y.numpy(), 8, ty)
        assert math.isclose(calculate_psnr(x.to(dev), y.to(dev), 8, test_y_channel=ty), ref, rel_tol=1e-9)


@pytest.mark.gpu
def test_gpu_ssim_matches_reference():
    """hatb200_ssim_u8 against the reference-generated values (basicsr/metrics/psnr_ssim.py:85-129) and, at HR size with
    ragged tiles, against the oracle; 1e-9 relative (float64 sums in a different order than cv2.filter2D)."""
    from super_resolution_b200 import calculate_ssim
    z, _ = _gold()
    ssim = json.loads(str(z["ssim"]))
    dev = "cuda:0"
    ia, ib = torch.from_numpy(z["a_bgr"]).to(dev), torch.from_numpy(z["b_bgr"]).to(dev)
    for cb in (0, 4):
        for ty in (False, True):
            assert calculate_ssim(ia, ib, cb, test_y_channel=ty) == pytest.approx(ssim[f"cb{cb}_y{int(ty)}"], rel=1e-9)
    assert calculate_ssim(ia, ia, 0) == pytest.approx(1.0, abs=1e-12)
    assert calculate_ssim(ia.permute(2, 0, 1), ib.permute(2, 0, 1), 4, input_order="CHW") == \
        pytest.approx(ssim["cb4_y0"], rel=1e-9)
    _, _, gray, _ = postproc_inputs()
    g2 = torch.from_numpy(IO.tensor2img((gray * 0.97).numpy())).to(dev)
    gr = torch.from_numpy(z["gray"]).to(dev)
    assert calculate_ssim(gr, g2, 2, test_y_channel=True) == pytest.approx(ssim["gray_cb2_y1"], rel=1e-9)
    assert calculate_ssim(gr, g2, 0) == pytest.approx(ssim["gray_cb0_y0"], rel=1e-9)
    # more tiles than blocks (grid-stride) and ragged edges, against the oracle
    g = torch.Generator().manual_seed(13)
    x = torch.randint(0, 256, (1083, 1531, 3), generator=g, dtype=torch.uint8)
    y = (x.int() + torch.randint(-9, 10, x.shape, generator=g)).clamp(0, 255).to(torch.uint8)
    for ty in (False, True):
        ref = IO.calculate_ssim(x.numpy(), y.numpy(), 4, ty)
        assert math.isclose(calculate_ssim(x.to(dev), y.to(dev), 4, test_y_channel=ty), ref, rel_tol=1e-9)
    with pytest.raises(RuntimeError):
        calculate_ssim(ia[:12, :12], ib[:12, :12], 2)          # 8x8 after the crop: smaller than the window
