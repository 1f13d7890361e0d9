"""CPU: pin the oracle restatement (oracle/hat_oracle.py) against fixtures produced by the
unmodified reference (oracle/make_golden.py), and against the live reference when mounted."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import hat_oracle as O
from oracle import load_ref
from tests.helpers import FORWARD_CASES, GOLDEN, load_case

SMALL = [c for c in FORWARD_CASES if c != "hat_full_x4_64x64"]


@pytest.mark.parametrize("name", SMALL)
def test_oracle_matches_reference_golden(name):
    cfg, sd, x, y_ref = load_case(name)
    y = O.hat_forward(x, sd, cfg)
    assert y.shape == y_ref.shape
    scale = max(1.0, y_ref.abs().max().item())
    assert (y - y_ref).abs().max().item() <= 1e-4 * scale


@pytest.mark.slow
def test_oracle_matches_reference_golden_full_hat():
    cfg, sd, x, y_ref = load_case("hat_full_x4_64x64")
    y = O.hat_forward(x, sd, cfg)
    assert (y - y_ref).abs().max().item() <= 1e-4 * max(1.0, y_ref.abs().max().item())


def test_index_tables_bit_exact():
    z = np.load(os.path.join(GOLDEN, "index_tables.npz"))
    assert np.array_equal(O.rpi_oca(16, 0.5), z["rpi_oca"])
    assert np.array_equal(O.rpi_sa(16), z["rpi_sa"])
    # negative entries wrap onto a bijection of 0..1520
    w = O.wrap_index(z["rpi_oca"], 39 * 39)
    assert z["rpi_oca"].min() == -880 and z["rpi_oca"].max() == 640
    assert w.min() == 0 and w.max() == 1520 and len(np.unique(w)) == 1521


def test_tile_rects_bit_exact():
    with open(os.path.join(GOLDEN, "tile_rects.json")) as f:
        gold = json.load(f)
    for key, rects in gold.items():
        hw, ts, pad, s = key.split("_")
        h, w = (int(v) for v in hw.split("x"))
        mine = O.tile_rects(h, w, int(ts[2:]), int(pad[3:]), int(s[1:]))
        assert mine == rects, key


def test_window_and_unfold_indices_match_torch():
    h, w, ws, ows = 32, 48, 16, 24
    idx = torch.arange(h * w, dtype=torch.float32).view(1, 1, h, w)
    # window_partition semantics (hat_arch.py:124-128)
    win = idx.view(1, h // ws, ws, w // ws, ws, 1).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws)
    assert np.array_equal(O.window_token_index(h, w, ws), win.numpy().astype(np.int64))
    # nn.Unfold(ows, stride ws, pad 4) semantics (hat_arch.py:296-297); +1 so that zero padding reads as -1
    unf = torch.nn.functional.unfold(idx + 1, kernel_size=ows, stride=ws, padding=(ows - ws) // 2)
    unf = unf[0].transpose(0, 1).numpy().astype(np.int64) - 1
    assert np.array_equal(O.overlap_window_token_index(h, w, ws, ows), unf)


def test_pixel_shuffle_map_matches_torch():
    for r in (2, 3):
        c, h, w = 2, 3, 5
        src = torch.arange(c * r * r * h * w, dtype=torch.float32).view(1, c * r * r, h, w)
        out = torch.nn.functional.pixel_shuffle(src, r)[0].numpy().astype(np.int64)
        assert np.array_equal(O.pixel_shuffle_source(c, h, w, r), out)


def test_batch_equals_per_sample():
    cfg, sd, x, _ = load_case("mini_x4_b3_32x32")
    y = O.hat_forward(x, sd, cfg)
    y1 = O.hat_forward(x[1:2], sd, cfg)
    assert (y[1:2] - y1).abs().max().item() <= 1e-5


def test_bad_size_raises():
    cfg, sd, x, _ = load_case("mini_x4_32x32")
    with pytest.raises(RuntimeError):
        O.hat_forward(torch.rand(1, 3, 24, 32), sd, cfg)


@pytest.mark.skipif(not load_ref.reference_available(), reason="reference not mounted (GPU box)")
def test_oracle_matches_live_reference():
    cfg, sd, x, y_gold = load_case("mini_x2_ocabesc")
    net = load_ref.build_reference(cfg)
    net.load_state_dict(sd, strict=True)
    y_ref = load_ref.reference_forward(net, x)
    assert torch.equal(y_ref, y_gold)
    assert (O.hat_forward(x, sd, cfg) - y_ref).abs().max().item() <= 1e-4
    # tiled path vs a hand loop over the reference net
    img = torch.rand(1, 3, 40, 56, generator=torch.Generator().manual_seed(3))
    cfg2, sd2, _, _ = load_case("mini_x2_32x48")
    net2 = load_ref.build_reference(cfg2)
    net2.load_state_dict(sd2, strict=True)
    a = O.tiled_forward(img, sd2, cfg2, 32, 16)
    b = O.tiled_forward(img, sd2, cfg2, 32, 16, forward=lambda t: load_ref.reference_forward(net2, t))
    assert (a - b).abs().max().item() <= 1e-4
