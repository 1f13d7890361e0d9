"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference's output post-processing, the step right after
the hot path.  Only tests/ may import this; the product path is the CUDA kernel
`hatb200_tensor2img` / `hatb200_psnr_u8` / `hatb200_ssim_u8`.  Pinned against the unmodified reference functions by
`oracle/make_golden.py::gen_postproc` -> tests/golden/postproc.npz (and live when /root/reference is mounted)."""
import numpy as np


def tensor2img(x, rgb2bgr=True, min_max=(0, 1)):
    """basicsr/utils/img_util.py:38-95 for one (1,C,H,W) / (C,H,W) / (H,W) float array, out_type uint8:
    squeeze(0), clamp, (x - min) / (max - min), CHW -> HWC, RGB -> BGR (cv2.cvtColor is a channel flip), * 255,
    numpy round (half to even), uint8.  (The reference also clamps its fp32 CPU argument in place, `.clamp_` on a view
    at img_util.py:67; neither this restatement nor the device kernel writes to its input.)"""
    x = np.asarray(x, dtype=np.float32)
    if x.ndim >= 1 and x.shape[0] == 1:
        x = x[0]
    x = np.clip(x, np.float32(min_max[0]), np.float32(min_max[1]))
    x = (x - np.float32(min_max[0])) / np.float32(min_max[1] - min_max[0])
    if x.ndim == 3:
        img = x.transpose(1, 2, 0)
        if img.shape[2] == 1:
            img = img[:, :, 0]
        elif rgb2bgr:
            img = img[:, :, ::-1]
    elif x.ndim == 2:
        img = x
    else:
        raise TypeError(f'Only support 4D, 3D or 2D tensor. But received with dimension: {x.ndim}')
    return (img * np.float32(255.0)).round().astype(np.uint8)


def to_y_channel(img):
    """basicsr/metrics/metric_util.py:32-45 with bgr2ycbcr(y_only=True) of basicsr/utils/color_util.py:60-68,
    :145-151, :179-183 inlined."""
    img = img.astype(np.float32) / 255.
    if img.ndim == 3 and img.shape[2] == 3:
        y = np.dot(img, [24.966, 128.553, 65.481]) + 16.0      # float64
        y /= 255.
        img = y.astype(np.float32)[..., None]
    return img * 255.


def calculate_psnr(img, img2, crop_border, test_y_channel=False):
    """basicsr/metrics/psnr_ssim.py:12-50 (input_order 'HWC')."""
    assert img.shape == img2.shape
    if img.ndim == 2:
        img, img2 = img[..., None], img2[..., None]
    if crop_border != 0:
        img = img[crop_border:-crop_border, crop_border:-crop_border, ...]
        img2 = img2[crop_border:-crop_border, crop_border:-crop_border, ...]
    if test_y_channel:
        img, img2 = to_y_channel(img), to_y_channel(img2)
    mse = np.mean((img.astype(np.float64) - img2.astype(np.float64)) ** 2)
    if mse == 0:
        return float('inf')
    return 10. * np.log10(255. * 255. / mse)


def gaussian_kernel_11():
    """cv2.getGaussianKernel(11, 1.5) (float64): exp(-x^2 / (2 sigma^2)) normalised to sum 1."""
    x = np.arange(11, dtype=np.float64) - 5.0
    k = np.exp(-0.5 / (1.5 * 1.5) * x * x)
    return k * (1.0 / k.sum())


def _ssim(img, img2):
    """basicsr/metrics/psnr_ssim.py:170-198 for one float64 channel: cv2.filter2D(...)[5:-5, 5:-5] is the 11x11
    correlation over the valid region."""
    c1, c2 = (0.01 * 255) ** 2, (0.03 * 255) ** 2
    k = gaussian_kernel_11()
    window = np.outer(k, k)
    h, w = img.shape

    def filt(z):
        out = np.zeros((h - 10, w - 10), dtype=np.float64)
        for i in range(11):
            for j in range(11):
                out += window[i, j] * z[i:i + h - 10, j:j + w - 10]
        return out

    mu1, mu2 = filt(img), filt(img2)
    mu1_sq, mu2_sq, mu1_mu2 = mu1 ** 2, mu2 ** 2, mu1 * mu2
    sigma1_sq = filt(img ** 2) - mu1_sq
    sigma2_sq = filt(img2 ** 2) - mu2_sq
    sigma12 = filt(img * img2) - mu1_mu2
    ssim_map = ((2 * mu1_mu2 + c1) * (2 * sigma12 + c2)) / ((mu1_sq + mu2_sq + c1) * (sigma1_sq + sigma2_sq + c2))
    return ssim_map.mean()


def calculate_ssim(img, img2, crop_border, test_y_channel=False):
    """basicsr/metrics/psnr_ssim.py:85-129 (input_order 'HWC')."""
    assert img.shape == img2.shape
    if img.ndim == 2:
        img, img2 = img[..., None], img2[..., None]
    if crop_border != 0:
        img = img[crop_border:-crop_border, crop_border:-crop_border, ...]
        img2 = img2[crop_border:-crop_border, crop_border:-crop_border, ...]
    if test_y_channel:
        img, img2 = to_y_channel(img), to_y_channel(img2)
    img, img2 = img.astype(np.float64), img2.astype(np.float64)
    return np.array([_ssim(img[..., i], img2[..., i]) for i in range(img.shape[2])]).mean()
