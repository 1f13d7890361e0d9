"""super_resolution_b200 -- B200-native forward of the reference `HAT` super-resolution network.

Public surface (mirrors the reference's own for this path):
  * `HAT`            -- drop-in for `hat.archs.hat_arch.HAT` (same ctor, same state_dict)
  * `HATModel`       -- HATModel.pre_process/process/tile_process/post_process on device, tiles optionally
                        sharded over ranks (reference: HAT/hat/models/hat_model.py:16-112)
  * `tensor2img`, `calculate_psnr`, `calculate_ssim` -- the post-processing right after the path (basicsr/utils/img_util.py:38-95,
                        basicsr/metrics/psnr_ssim.py:12-50) as device kernels
  * `load_network`, `save_network` -- the reference's checkpoint file format (basicsr/models/base_model.py:207-315)
The arithmetic lives in `libhatb200.so` (csrc/, include/hatb200.h).
"""
from .hat_arch import HAT, HATX  # noqa: F401
from .hat_model import HATModel, calculate_psnr, calculate_ssim, load_network, save_network, tensor2img  # noqa: F401

__all__ = ["HAT", "HATX", "HATModel", "tensor2img", "calculate_psnr", "calculate_ssim", "load_network", "save_network"]
__version__ = "0.1.0"
