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
    from super_resolution_b200.hat_arch import HAT
    net = HAT(**cfg)
    sd = O.randomized_state_dict(net.state_dict(), int(z["sd_seed"]))
    shape = tuple(int(v) for v in z["x_shape"])
    x = torch.rand(shape, generator=torch.Generator().manual_seed(int(z["x_seed"])), dtype=torch.float32)
    return cfg, sd, x, torch.from_numpy(z["y"])


def psnr(a, b, peak=1.0):
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)
