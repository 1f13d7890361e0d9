"""CPU: the drop-in boundary -- state_dict layout, constructor semantics, C-ABI exports."""
import copy
import os
import re
import subprocess

import numpy as np
import pytest
import torch

from oracle import hat_oracle as O
from super_resolution_b200 import HAT, _lib
from tests.helpers import GOLDEN, load_manifest
import json

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", ["HAT-S_x4", "HAT_x2", "HAT_x3", "HAT_x4", "HAT-L_x4", "mini_x2_ocabesc"])
def test_state_dict_layout_matches_reference(name):
    e = load_manifest()[name]
    net = HAT(**e["cfg"])
    sd = net.state_dict()
    assert list(sd.keys()) == list(e["keys"].keys())          # same keys, same order
    for k, (shape, dtype) in e["keys"].items():
        assert list(sd[k].shape) == shape and str(sd[k].dtype) == dtype, k
    assert sum(p.numel() for p in net.parameters()) == e["params"]


def test_strict_load_roundtrip_and_index_buffers():
    e = load_manifest()["HAT_x4"]
    net = HAT(**e["cfg"])
    sd = O.randomized_state_dict(net.state_dict(), 3)
    net.load_state_dict(sd, strict=True)
    for k, v in net.state_dict().items():
        assert torch.equal(v, sd[k])
    z = np.load(os.path.join(GOLDEN, "index_tables.npz"))
    assert np.array_equal(net.relative_position_index_OCA.numpy(), z["rpi_oca"])
    assert np.array_equal(net.relative_position_index_SA.numpy(), z["rpi_sa"])
    m = HAT(**dict(e["cfg"], embed_dim=48, depths=[1], num_heads=[6]))
    assert np.array_equal(m.calculate_mask((32, 48)).numpy(), z["mask_32x48"])


def test_constructor_semantics():
    net = HAT(upscale=4, window_size=16, embed_dim=48, depths=[1], num_heads=[6], upsampler='pixelshuffle',
              squeeze_factor=99, esc_use_dynamic=False, use_checkpoint=True, some_unknown_kwarg=1)
    assert net.no_weight_decay() == {'absolute_pos_embed'}
    assert net.no_weight_decay_keywords() == {'relative_position_bias_table'}
    assert net.shift_size == 8 and net.num_layers == 1 and net.num_features == 48
    assert net.mean.shape == (1, 3, 1, 1)
    assert HAT(in_chans=1, window_size=16, embed_dim=48, depths=[1], num_heads=[6]).mean.shape == (1, 1, 1, 1)
    with pytest.raises(ValueError):
        HAT(resi_connection='3conv', window_size=16, embed_dim=48, depths=[1], num_heads=[6])
    with pytest.raises(ValueError):
        HAT(upscale=5, upsampler='pixelshuffle', window_size=16, embed_dim=48, depths=[1], num_heads=[6])
    with pytest.raises(NotImplementedError):
        HAT(ape=True, window_size=16, embed_dim=48, depths=[1], num_heads=[6])
    # fresh init follows the reference: zero dynamic-kernel head, identity LayerNorm, zero Linear bias
    blk = net.layers[0].residual_group.blocks[0]
    assert blk.esc_attn.core.plk.dwc_proj[3].weight.abs().max() == 0
    assert torch.equal(blk.norm1.weight, torch.ones(48)) and blk.mlp.fc1.bias.abs().max() == 0
    assert copy.deepcopy(net).state_dict().keys() == net.state_dict().keys()


def test_no_cpu_path():
    net = HAT(upscale=2, window_size=16, embed_dim=48, depths=[1], num_heads=[6], upsampler='pixelshuffle').eval()
    with pytest.raises(RuntimeError, match="no CPU path"):
        net(torch.rand(1, 3, 16, 16))
    with pytest.raises(RuntimeError, match="inference-only"):
        net.train()(torch.rand(1, 3, 16, 16))


def test_library_loads_and_exports_every_declared_symbol():
    l = _lib.lib()
    assert l.hatb200_version() == 101
    header = open(os.path.join(ROOT, "include", "hatb200.h")).read()
    product, eng = re.subn(r"#ifdef HATB200_ENGINEERING.*?#endif", "", header, flags=re.S)
    assert eng == 1
    declared = sorted(set(re.findall(r"\b(hatb200_[a-z0-9_]+)\s*\(", product)))
    assert declared == _lib.exported_symbols()
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    for sym in declared:
        assert re.search(rf"\bT {sym}\b", nm), f"{sym} not exported by libhatb200.so"
    # the engineering probes are declared under HATB200_ENGINEERING and are not part of the product library
    exported = set(re.findall(r"\bT (hatb200_[a-z0-9_]+)\b", nm))
    assert exported == set(declared), exported ^ set(declared)
    assert __import__("ctypes").sizeof(_lib.Config) == 372   # sizeof(hatb200_config); create() re-checks it


def test_tile_plan_c_abi_bit_exact():
    with open(os.path.join(GOLDEN, "tile_rects.json")) as f:
        gold = json.load(f)
    for key, rects in gold.items():
        hw, ts, pad, s = key.split("_")
        h, w = (int(v) for v in hw.split("x"))
        assert _lib.tile_plan(h, w, int(ts[2:]), int(pad[3:]), int(s[1:])) == rects, key


def test_create_validates_without_gpu_compute():
    import ctypes as C
    l = _lib.lib()
    cfg = _lib.Config()
    cfg.struct_size = 3
    h = C.c_void_p()
    assert l.hatb200_create(C.byref(cfg), C.byref(h)) == -1
    assert b"ABI mismatch" in l.hatb200_last_error()
    # a well-formed config passes the ABI guard: without a GPU the only failure is "no CUDA device"
    from super_resolution_b200.engine import _config_from_module
    net = HAT(upscale=2, window_size=16, embed_dim=48, depths=[1], num_heads=[6], upsampler='pixelshuffle')
    cfg = _config_from_module(net)
    rc = l.hatb200_create(C.byref(cfg), C.byref(h))
    if torch.cuda.is_available():
        assert rc == 0
        assert l.hatb200_num_params(h) == len(net.state_dict())
        keys = [l.hatb200_param_key(h, i).decode() for i in range(l.hatb200_num_params(h))]
        assert keys == list(net.state_dict().keys())
        l.hatb200_destroy(h)
    else:
        assert rc == -2 and b"no CPU path" in l.hatb200_last_error()
    cfg.upscale = 5
    assert l.hatb200_create(C.byref(cfg), C.byref(h)) == -1 and b"not supported" in l.hatb200_last_error()
