"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference `HAT` class.

Only usable in the build container where `/root/reference` is mounted (it does
not exist on the GPU box).  Used by `oracle/make_golden.py` to generate the
committed fixtures under `tests/golden/` and by the CPU tests that pin the
restatement in `oracle/hat_oracle.py` against the live reference.

The reference package `hat` cannot be imported normally: `hat/archs/__init__.py`
imports every `*_arch.py` and `hitsrf_arch.py` needs `timm` (absent).  We insert
empty `hat` / `hat.archs` packages and load `esc_arch.py` + `hat_arch.py` by path
(SURVEY.md section 8c).  Nothing is copied; the reference stays read-only.
"""
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("HAT_REFERENCE_ROOT", "/root/reference")
_ARCH_DIR = os.path.join(REF_ROOT, "HAT", "hat", "archs")
_BASICSR_PARENT = os.path.join(REF_ROOT, "HAT", "ESC")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(_ARCH_DIR, "hat_arch.py"))


def load_reference_hat():
    """Return the reference module `hat.archs.hat_arch` (class `HAT` inside)."""
    if not reference_available():
        raise FileNotFoundError(f"reference not mounted at {REF_ROOT}")
    if "hat.archs.hat_arch" in sys.modules:
        return sys.modules["hat.archs.hat_arch"]
    sys.dont_write_bytecode = True
    if _BASICSR_PARENT not in sys.path:
        sys.path.insert(0, _BASICSR_PARENT)
    for name, path in (("hat", os.path.join(REF_ROOT, "HAT", "hat")), ("hat.archs", _ARCH_DIR)):
        if name not in sys.modules:
            pkg = types.ModuleType(name)
            pkg.__path__ = [path]
            sys.modules[name] = pkg
    mods = {}
    for short in ("esc_arch", "hat_arch"):
        full = f"hat.archs.{short}"
        spec = importlib.util.spec_from_file_location(full, os.path.join(_ARCH_DIR, f"{short}.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        mods[short] = mod
    return mods["hat_arch"]


def load_reference_hatx():
    """Return the reference module `hat.archs.hatx_arch` (class `HATX`, the sibling variant of SURVEY.md 8(f4))."""
    load_reference_hat()
    full = "hat.archs.hatx_arch"
    if full in sys.modules:
        return sys.modules[full]
    spec = importlib.util.spec_from_file_location(full, os.path.join(_ARCH_DIR, "hatx_arch.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[full] = mod
    spec.loader.exec_module(mod)
    return mod


def build_reference_hatx(cfg: dict):
    mod = load_reference_hatx()
    net = mod.HATX(**{k: v for k, v in cfg.items() if k != "arch"})
    net.eval()
    return net


def build_reference(cfg: dict):
    """Instantiate the reference `HAT(**cfg)` in eval mode."""
    mod = load_reference_hat()
    net = mod.HAT(**cfg)
    net.eval()
    return net


def reference_forward(net, x):
    """Reference eval forward; the reference's eval branch only supports B=1
    (esc_arch.py:121), so a batch is a per-sample loop."""
    import torch
    with torch.no_grad():
        return torch.cat([net(x[i:i + 1]) for i in range(x.shape[0])], dim=0)
