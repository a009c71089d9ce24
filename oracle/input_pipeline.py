"""ORACLE (test infrastructure, not product): CPU restatement of the reference's input transform.

Reference: data_and_toy_model.py:13-36 — transforms.Resize(224) -> RandomHorizontalFlip -> ToTensor ->
Normalize((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)) applied to CIFAR-10's 32x32 RGB uint8 images
(test: the same without the flip). The arithmetic lives in third-party code that is not under /root/reference:
  * torchvision 0.26.0 (unpinned by the reference, requirements.txt:2) transforms/functional.py resize ->
    _functional_pil.resize -> PIL.Image.resize(size, BILINEAR), to_tensor (uint8 HWC -> float CHW / 255),
    normalize ((x - mean) / std);
  * Pillow 12.2.0 src/libImaging/Resample.c: ImagingResample = horizontal pass then vertical pass, each
    with per-output-pixel normalised filter coefficients in 22-bit fixed point
    (precompute_coeffs + normalize_coeffs_8bpc + ImagingResampleHorizontal_8bpc / Vertical_8bpc), rounding to
    uint8 BETWEEN the two passes.
Pinned: oracle/make_golden.py runs the real torchvision transform here (Pillow present in the build container) and
tests/golden/input_pipeline.npz holds its outputs; tests/test_oracle_golden.py checks this file against them
bit for bit (uint8 stage) and to 1 ulp (float stage).
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2
MEAN = (0.4914, 0.4822, 0.4465)
STD = (0.2023, 0.1994, 0.2010)


def _bilinear(x: float) -> float:
    if x < 0.0:
        x = -x
    return 1.0 - x if x < 1.0 else 0.0


def resample_coeffs(in_size: int, out_size: int):
    """(xmin[out], count[out], coef[out][ksize] int32) of Resample.c precompute_coeffs + normalize_coeffs_8bpc
    for the bilinear filter (support 1.0) over the whole axis (box = [0, in_size))."""
    scale = filterscale = in_size / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    xmins = np.zeros(out_size, np.int32)
    counts = np.zeros(out_size, np.int32)
    coefs = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        k = [_bilinear((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for w in k:
            ww += w
        if ww != 0.0:
            k = [w / ww for w in k]
        for x, w in enumerate(k):
            coefs[xx, x] = int(-0.5 + w * (1 << PRECISION_BITS)) if w < 0 else int(0.5 + w * (1 << PRECISION_BITS))
        xmins[xx] = xmin
        counts[xx] = xmax
    return xmins, counts, coefs


def _pass(img: np.ndarray, axis: int, out_size: int) -> np.ndarray:
    """One 8-bit resample pass along `axis` of a [H][W][C] uint8 image."""
    xmins, counts, coefs = resample_coeffs(img.shape[axis], out_size)
    src = np.moveaxis(img, axis, 0).astype(np.int64)
    out = np.empty((out_size,) + src.shape[1:], np.uint8)
    for xx in range(out_size):
        acc = np.full(src.shape[1:], 1 << (PRECISION_BITS - 1), np.int64)
        for x in range(counts[xx]):
            acc += src[xmins[xx] + x] * int(coefs[xx, x])
        out[xx] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def resize_bilinear_u8(img_hwc: np.ndarray, size: int = 224) -> np.ndarray:
    """PIL.Image.resize((size, size), BILINEAR) of a [H][W][C] uint8 image: horizontal pass, then vertical."""
    return _pass(_pass(img_hwc, 1, size), 0, size)


def transform(img_hwc: np.ndarray, flip: bool, size: int = 224) -> np.ndarray:
    """Resize(size) -> (flip) -> ToTensor -> Normalize: [H][W][3] uint8 -> [3][size][size] float32."""
    r = resize_bilinear_u8(img_hwc, size)
    if flip:
        r = r[:, ::-1, :]
    t = np.ascontiguousarray(r.transpose(2, 0, 1)).astype(np.float32) / np.float32(255.0)     # to_tensor: .div(255)
    mean = np.asarray(MEAN, np.float32).reshape(3, 1, 1)
    std = np.asarray(STD, np.float32).reshape(3, 1, 1)
    return ((t - mean) / std).astype(np.float32)                                            # normalize: sub_().div_()


def transform_batch(imgs_bhwc: np.ndarray, flips, size: int = 224) -> np.ndarray:
    return np.stack([transform(imgs_bhwc[i], bool(flips[i]), size) for i in range(imgs_bhwc.shape[0])])
