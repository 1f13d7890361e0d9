"""The drop-in `HATModel(opt)` (reference HAT/hat/models/hat_model.py:13-185): construction from the option dict as
`basicsr.models.build_model` does it, reference-layout checkpoint files, `nondist_validation` with device-side
tensor2img / PSNR / SSIM, and registration in basicsr's registries when basicsr is importable."""
import os
import subprocess
import sys
import types

import numpy as np
import pytest
import torch

from oracle import hat_oracle as O
from oracle import img_oracle as IO
from oracle import load_ref
from tests.helpers import load_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not load_ref.reference_available(), reason="reference (vendored basicsr) not mounted")
def test_registered_in_basicsr_registries_under_the_reference_names():
    """With the reference's vendored basicsr on the path and this package imported INSTEAD of hat.archs / hat.models,
    build_network / build_model resolve 'HAT' / 'HATModel' to the B200 classes (hat_arch.py:607, hat_model.py:13)."""
    code = ("import sys; sys.dont_write_bytecode = True; sys.path.insert(0, %r); sys.path.insert(0, %r);"
            "import super_resolution_b200 as P;"
            "from basicsr.utils.registry import ARCH_REGISTRY, MODEL_REGISTRY;"
            "assert ARCH_REGISTRY.get('HAT') is P.HAT; assert MODEL_REGISTRY.get('HATModel') is P.HATModel; print('ok')"
            % (os.path.join(load_ref.REF_ROOT, "HAT", "ESC"), ROOT))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]


def test_hatmodel_opt_is_inference_only_and_needs_cuda():
    from super_resolution_b200 import HATModel
    with pytest.raises(NotImplementedError):
        HATModel({"is_train": True, "network_g": {"type": "HAT"}})
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            HATModel({"is_train": False, "network_g": {"type": "HAT", "window_size": 16}})


def _opt(cfg, **extra):
    opt = {"name": "b200", "model_type": "HATModel", "scale": cfg["upscale"], "num_gpu": 1, "is_train": False,
           "dist": False, "rank": 0, "network_g": dict(cfg, type="HAT"), "path": {}, "val": {"suffix": None}}
    opt.update(extra)
    return opt


@pytest.mark.gpu
def test_reference_layout_checkpoint_loads_through_opt_and_matches_golden(tmp_path):
    """f3: a `.pth` as the reference's save_network writes it from a DDP-wrapped EMA model (`params_ema`, `module.`
    prefixes; base_model.py:207-252) -> HATModel(opt) -> forward == the reference-generated golden."""
    from super_resolution_b200 import HATModel
    cfg, sd, x, y_ref = load_case("mini_x4_32x32")
    ck = tmp_path / "net_g_latest.pth"
    torch.save({"params": {("module." + k): torch.zeros_like(v) for k, v in sd.items()},
                "params_ema": {("module." + k): v for k, v in sd.items()}}, ck)
    opt = _opt(cfg)
    opt["path"] = {"pretrain_network_g": str(ck), "param_key_g": "params_ema", "strict_load_g": True}
    model = HATModel(opt)
    assert next(model.net_g.parameters()).is_cuda and not model.net_g.training
    y = model.test(x.to("cuda:0"))
    err = (y.cpu() - y_ref).abs().max().item()
    assert err <= 1e-4 * max(1.0, y_ref.abs().max().item()), err
    # a file without params_ema falls back to params (base_model.py:304-306)
    torch.save({"params": sd}, ck)
    model2 = HATModel(opt)
    assert torch.equal(model2.test(x.to("cuda:0")), y)


@pytest.mark.gpu
@pytest.mark.parametrize("tile", [None, (32, 16)])
def test_nondist_validation_matches_the_reference_pipeline(tmp_path, tile):
    """hat_model.py:114-185 image by image; metrics and the saved image against the oracle pipeline
    (tiled_forward -> tensor2img -> calculate_psnr / calculate_ssim)."""
    from super_resolution_b200 import HATModel
    cfg, sd, _, _ = load_case("mini_x2_32x48")
    mean = torch.tensor([0.4488, 0.4371, 0.4040]).view(1, 3, 1, 1)
    probe = torch.rand(1, 3, 32, 32, generator=torch.Generator().manual_seed(1))
    sd = O.scale_tail(sd, 0.4 / (O.hat_forward(probe, sd, cfg) - mean).abs().max().item())
    ck = tmp_path / "w.pth"
    torch.save({"params": sd}, ck)
    g = torch.Generator().manual_seed(2)
    data = []
    for i, (h, w) in enumerate([(37, 45), (48, 64)]):
        lq = torch.rand(1, 3, h, w, generator=g)
        data.append({"lq": lq, "gt": torch.rand(1, 3, 2 * h, 2 * w, generator=g), "lq_path": [f"/data/img{i}.png"]})
    loader = types.SimpleNamespace(dataset=types.SimpleNamespace(opt={"name": "Set2"}))
    loader = type("L", (list,), {})(data)
    loader.dataset = types.SimpleNamespace(opt={"name": "Set2"})
    opt = _opt(cfg)
    opt["path"] = {"pretrain_network_g": str(ck), "visualization": str(tmp_path / "vis")}
    opt["val"] = {"suffix": None, "metrics": {
        "psnr": {"type": "calculate_psnr", "crop_border": 2, "test_y_channel": True},
        "ssim": {"type": "calculate_ssim", "crop_border": 2, "test_y_channel": True, "better": "higher"}}}
    if tile:
        opt["tile"] = {"tile_size": tile[0], "tile_pad": tile[1]}
    model = HATModel(opt)
    model.validation(loader, current_iter=7, tb_logger=None, save_img=True)
    want = {"psnr": 0.0, "ssim": 0.0}
    import cv2
    for i, d in enumerate(data):
        if tile:
            y = O.tiled_forward(d["lq"], sd, cfg, tile[0], tile[1])
        else:
            ph, pw = O.mod_pad(d["lq"].shape[2], d["lq"].shape[3], 16)
            y = O.hat_forward(torch.nn.functional.pad(d["lq"], (0, pw, 0, ph), "reflect"), sd, cfg)
            y = y[:, :, :2 * d["lq"].shape[2], :2 * d["lq"].shape[3]]
        sr, gt = IO.tensor2img(y.numpy()), IO.tensor2img(d["gt"].numpy())
        want["psnr"] += IO.calculate_psnr(sr, gt, 2, True) / len(data)
        want["ssim"] += IO.calculate_ssim(sr, gt, 2, True) / len(data)
        saved = cv2.imread(str(tmp_path / "vis" / "Set2" / f"img{i}_b200.png"), cv2.IMREAD_UNCHANGED)
        assert saved.shape == sr.shape
        diff = np.abs(saved.astype(np.int32) - sr.astype(np.int32))
        assert diff.max() <= 1 and (diff != 0).mean() < 1e-3        # fp32 summation order at exact .5 ties only
    assert model.metric_results["psnr"] == pytest.approx(want["psnr"], rel=1e-4)
    assert model.metric_results["ssim"] == pytest.approx(want["ssim"], rel=1e-4)
    assert model.best_metric_results["Set2"]["psnr"]["iter"] == 7


@pytest.mark.gpu
def test_packed_weight_cache_keyed_by_state_dict_hash(tmp_path):
    """f3: the packed (bf16 / tcgen05-layout) weight copies as a blob on disk keyed by the state_dict hash; a hit
    rebuilds the handle without set_param / re-layout kernels and the forward is bit-identical; other weights miss."""
    from super_resolution_b200 import HAT
    cfg, sd, x, y_ref = load_case("mini_x4_32x32")
    xd = x.to("cuda:0")

    def net_with(sd_, cache):
        net = HAT(**cfg).eval()
        net.load_state_dict(sd_, strict=True)
        net = net.to("cuda:0")
        net.engine().pack_cache_dir = str(cache) if cache else None
        return net

    base = net_with(sd, None)
    y0, y0b = base(xd), base(xd.bfloat16())
    first = net_with(sd, tmp_path)
    assert torch.equal(first(xd), y0) and first.engine().pack_cache_hit is False
    files = list(tmp_path.glob("*.hbpk"))
    assert len(files) == 1 and files[0].stat().st_size > 100000
    second = net_with(sd, tmp_path)
    assert torch.equal(second(xd), y0) and second.engine().pack_cache_hit is True
    assert torch.equal(second(xd.bfloat16()), y0b)
    sd2 = {k: (v * 1.01 if k == "conv_last.weight" else v) for k, v in sd.items()}
    third = net_with(sd2, tmp_path)
    assert not torch.equal(third(xd), y0) and third.engine().pack_cache_hit is False
    assert len(list(tmp_path.glob("*.hbpk"))) == 2
    # a truncated / foreign blob is rejected by the library and silently repacked
    files[0].write_bytes(files[0].read_bytes()[:5000])
    fourth = net_with(sd, tmp_path)
    assert torch.equal(fourth(xd), y0) and fourth.engine().pack_cache_hit is False
