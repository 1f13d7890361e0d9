"""Shared helpers for the test-suite (golden loading, seeded weights/inputs)."""
import json
import os

import numpy as np
import torch

from oracle import hat_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

FORWARD_CASES = [
    "mini_x4_32x32", "mini_x2_32x48", "mini_x3_48x32", "mini_x4_16x16", "mini_x4_b3_32x32",
    "mini_x2_ocabesc", "mini_x2_identity", "mini_x4_range255", "hats_dims_x4_32x32",
    "hat_dims_x4_32x32", "hat_full_x4_64x64",
]


HATX_CASES = ["hatx_mini_x2_32x32", "hatx_mini_x2_topk", "hatx_mini_x2_yaml_focus", "hatx_dims_x2_32x32"]


def load_manifest():
    with open(os.path.join(GOLDEN, "state_dict_manifest.json")) as f:
        return json.load(f)


def template_from_keys(keys: dict, cfg: dict):
    """Zero state_dict with the reference's key order / shapes; integer buffers are
    filled from the oracle's index functions."""
    sd = {}
    for k, (shape, dtype) in keys.items():
        if dtype == "torch.int64":
            if k.endswith("_SA"):
                sd[k] = torch.from_numpy(O.rpi_sa(cfg["window_size"]))
            else:
                sd[k] = torch.from_numpy(O.rpi_oca(cfg["window_size"], cfg["overlap_ratio"]))
        else:
            sd[k] = torch.zeros(shape, dtype=torch.float32)
    return sd


def load_case(name):
    """-> (cfg, state_dict, x, y_reference) for a committed golden case.  The
    state_dict template comes from OUR module (the reference is not on the GPU
    box); key order equality with the reference is asserted by the manifest test."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    cfg = json.loads(str(z["cfg"]))
    net = build_module(cfg)
    sd = O.randomized_state_dict(net.state_dict(), int(z["sd_seed"]))
    shape = tuple(int(v) for v in z["x_shape"])
    x = torch.rand(shape, generator=torch.Generator().manual_seed(int(z["x_seed"])), dtype=torch.float32)
    return cfg, sd, x, torch.from_numpy(z["y"])


def build_module(cfg):
    """Our drop-in module for a golden's cfg: `HATX` when the cfg carries arch == 'HATX' (the key is the oracle's
    switch, not a constructor argument), else `HAT`."""
    from super_resolution_b200.hat_arch import HAT, HATX
    if cfg.get("arch") == "HATX":
        return HATX(**{k: v for k, v in cfg.items() if k != "arch"})
    return HAT(**cfg)


# round 2: full-depth variants at BASELINE sizes with image-range outputs -- the north_star's bars asserted literally
R2_CASES = [
    "r2_hat_x4_b16_64x64", "r2_hatl_x4_64x64", "r2_hats_x4_64x64", "r2_hat_x2_64x64", "r2_hat_x3_64x64",
    "r2_hat_x2_256x256", "r2_hat_x3_256x256", "r2_hat_x4_fresh_64x64", "r2_hatl_x4_fresh_64x64",
    "r2_hat_x4_tiled_540x960",
]


def load_case_r2(name):
    """-> dict(cfg, sd, x, y, tile, sub, y_shape, y_stats) for a round-2 golden (oracle/make_golden.py CASES_R2).
    `y` is the reference output, stored every sub[0]-th row / sub[1]-th column when the full image is too big to
    commit; y_stats = float64 (sum, sum of squares) of the FULL reference output."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    cfg = json.loads(str(z["cfg"]))
    from super_resolution_b200.hat_arch import HAT
    if str(z["weights"]) == "fresh":
        sd = O.fresh_state_dict(HAT, cfg, int(z["sd_seed"]))
        got, want = O.state_dict_checksum(sd), float(z["sd_checksum"])
        assert abs(got - want) <= 1e-9 * abs(want), \
            f"{name}: seeded fresh init differs from the one the golden was made with (torch {z['torch_version']})"
    else:
        sd = O.randomized_state_dict(HAT(**cfg).state_dict(), int(z["sd_seed"]))
        sd = O.scale_tail(sd, float(z["tail_scale"]))
    shape = tuple(int(v) for v in z["x_shape"])
    x = torch.rand(shape, generator=torch.Generator().manual_seed(int(z["x_seed"])), dtype=torch.float32)
    tile = tuple(int(v) for v in z["tile"])
    return dict(cfg=cfg, sd=sd, x=x, y=torch.from_numpy(z["y"]), tile=tile if tile[0] else None,
                sub=tuple(int(v) for v in z["sub"]), y_shape=tuple(int(v) for v in z["y_shape"]),
                y_stats=z["y_stats"])


def psnr(a, b, peak=1.0):
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)
