"""Parity at the north_star's LITERAL bars, on full-depth variants at BASELINE sizes, against goldens produced by
the unmodified reference (oracle/make_golden.py CASES_R2): fp32 path max-abs <= 1e-4, bf16 path PSNR >= 50 dB at
peak 1.0.  Outputs are image-range (fresh init, or randomised weights with conv_last scaled by the committed
`tail_scale`), so no |y|max normalisation is involved.  The measured error is in every assert message.

CPU half: the oracle restatement against the same goldens (bounded: the 64x64 cases and two samples of the batch).
"""
import numpy as np
import pytest
import torch

from oracle import hat_oracle as O
from tests.helpers import R2_CASES, load_case_r2, psnr

DEV = "cuda:0"
FP32_MAX_ABS = 1e-4      # north_star, literal
BF16_PSNR_DB = 50.0      # north_star, literal, peak = 1.0


def _run(case, dtype):
    from super_resolution_b200 import HAT, HATModel
    cfg = case["cfg"]
    net = HAT(**cfg).eval()
    net.load_state_dict(case["sd"], strict=True)
    net = net.to(DEV)
    x = case["x"].to(DEV).to(dtype)
    ws = cfg["window_size"]
    if case["tile"] or x.shape[2] % ws or x.shape[3] % ws:
        opt = {"network_g": {"window_size": ws}, "scale": cfg["upscale"]}
        if case["tile"]:
            opt["tile"] = {"tile_size": case["tile"][0], "tile_pad": case["tile"][1]}
        y = HATModel(net, opt).test(x)
    else:
        y = net(x)
    assert y.dtype == dtype and tuple(y.shape) == case["y_shape"]
    y = y.float().cpu()
    stats = np.array([y.double().sum().item(), (y.double() ** 2).sum().item()])
    return y[:, :, ::case["sub"][0], ::case["sub"][1]], stats


@pytest.mark.gpu
@pytest.mark.parametrize("name", R2_CASES)
def test_fp32_max_abs_1e4_vs_reference(name):
    case = load_case_r2(name)
    assert -0.05 <= case["y"].min().item() and case["y"].max().item() <= 1.05    # an image: peak 1.0 is the right peak
    y, stats = _run(case, torch.float32)
    err = (y - case["y"]).abs().max().item()
    assert err <= FP32_MAX_ABS, f"{name}: fp32 max-abs {err:.3e} > {FP32_MAX_ABS}"
    # the full image (not only the stored subsample) through its float64 moments
    rel = np.abs(stats - case["y_stats"]) / np.abs(case["y_stats"])
    assert rel.max() <= 1e-5, f"{name}: full-image moments differ by {rel}"
    print(f"{name}: fp32 max-abs {err:.3e}")


@pytest.mark.gpu
@pytest.mark.parametrize("name", R2_CASES)
def test_bf16_psnr_50db_at_peak_1_vs_reference(name):
    case = load_case_r2(name)
    y, _ = _run(case, torch.bfloat16)
    db = psnr(y, case["y"], peak=1.0)
    assert db >= BF16_PSNR_DB, f"{name}: bf16 PSNR {db:.2f} dB < {BF16_PSNR_DB} at peak 1.0"
    print(f"{name}: bf16 PSNR {db:.2f} dB at peak 1.0, max-abs {(y - case['y']).abs().max().item():.3e}")


CPU_CASES = ["r2_hats_x4_64x64", "r2_hat_x2_64x64", "r2_hat_x3_64x64", "r2_hat_x4_fresh_64x64", "r2_hatl_x4_64x64"]


@pytest.mark.parametrize("name", CPU_CASES)
def test_oracle_matches_r2_golden(name):
    case = load_case_r2(name)
    x = case["x"][:1]
    y = O.hat_forward(x, case["sd"], case["cfg"])
    err = (y - case["y"][:1]).abs().max().item()
    assert err <= FP32_MAX_ABS, f"{name}: oracle vs reference max-abs {err:.3e}"


def test_oracle_matches_r2_batch_golden_samples():
    case = load_case_r2("r2_hat_x4_b16_64x64")
    for i in (0, 15):
        y = O.hat_forward(case["x"][i:i + 1], case["sd"], case["cfg"])
        err = (y - case["y"][i:i + 1]).abs().max().item()
        assert err <= FP32_MAX_ABS, f"sample {i}: {err:.3e}"
