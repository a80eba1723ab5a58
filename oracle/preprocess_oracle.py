"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the reference's image pre-processing.

Reference call chain: `ImageProcessor.process` (`/root/reference/src/multimeditron/model/modalities/
image_modality.py:96`) -> `AutoImageProcessor` = transformers 5.5.0 `CLIPImageProcessor(TorchvisionBackend)`
(`HF:models/clip/image_processing_clip.py:25-44`) -> `TorchvisionBackend._preprocess`
(`HF:image_processing_backends.py:372-420`):

    pil_to_tensor (uint8 CHW)                                         image_processing_backends.py:127-128
    resize: shortest edge -> S, long edge int(S*long/short)           image_processing_backends.py:231-237,
            tvF.resize(uint8, BICUBIC, antialias=True)                 image_transforms.py:295-299, :251
    center_crop to (S, S)                                             image_processing_backends.py:338-362
    (x.float() - mean*255) / (std*255)   fused rescale+normalise      image_processing_backends.py:292-330

The uint8 antialiased bicubic resize is ATen's separable integer kernel (`aten/src/ATen/native/cpu/UpSampleKernel.cpp`,
`_compute_index_ranges_int16_weights` + `upsample_avx_bilinear_bicubic_uint8`; torch 2.11, not vendored in
/root/reference): per axis, double-precision weights of the a=-0.5 cubic with support stretched by the scale,
normalised, quantised to int16 at the largest precision that keeps the biggest weight below 2^15; horizontal pass
first, uint8 intermediate, then the vertical pass; each pass accumulates int32 from 2^(prec-1) and shifts.
Pinned bit-exact against torch itself (`tests/test_preprocess_oracle.py`, which runs torch's kernel on the host).
"""
from __future__ import annotations

import math

import numpy as np

OPENAI_CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)
OPENAI_CLIP_STD = (0.26862954, 0.26130258, 0.27577711)


def resize_output_size(h: int, w: int, shortest_edge: int) -> tuple[int, int]:
    """HF `get_resize_output_image_size(default_to_square=False)` (image_transforms.py:295-299)."""
    short, long = (w, h) if w <= h else (h, w)
    new_short, new_long = shortest_edge, int(shortest_edge * long / short)
    return (new_long, new_short) if w <= h else (new_short, new_long)


def center_crop_origin(h: int, w: int, ch: int, cw: int) -> tuple[int, int]:
    """image_processing_backends.py:360-361 (no padding branch: callers guarantee ch<=h, cw<=w)."""
    return int((h - ch) / 2.0), int((w - cw) / 2.0)


def _cubic_aa(x: float, a: float = -0.5) -> float:
    x = abs(x)
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1.0
    if x < 2.0:
        return (((x - 5.0) * x + 8.0) * x - 4.0) * a
    return 0.0


def aa_weights_int16(in_size: int, out_size: int):
    """Per output index: (xmin, xsize, int16 weights[ksize]) and the shift `prec` — ATen
    `_compute_indices_min_size_weights_aa` + `_compute_index_ranges_int16_weights`."""
    scale = in_size / out_size
    support = 2.0 * scale if scale >= 1.0 else 2.0
    ksize = int(math.ceil(np.float32(support))) * 2 + 1
    invscale = 1.0 / scale if scale >= 1.0 else 1.0
    xmin = np.zeros(out_size, np.int64)
    xsize = np.zeros(out_size, np.int64)
    wf = np.zeros((out_size, ksize), np.float64)
    wt_max = 0.0
    for i in range(out_size):
        center = scale * (i + 0.5)
        lo = max(int(center - support + 0.5), 0)
        n = min(int(center + support + 0.5), in_size) - lo
        n = max(0, min(n, ksize))
        tot = 0.0
        for j in range(n):
            w = _cubic_aa((j + lo - center + 0.5) * invscale)
            wf[i, j] = w
            tot += w
        if tot != 0.0:
            for j in range(n):
                wf[i, j] /= tot
        if n:
            wt_max = max(wt_max, float(wf[i, :n].max()))
        xmin[i], xsize[i] = lo, n
    prec = 0
    while prec < 22:
        if int(0.5 + wt_max * (1 << (prec + 1))) >= (1 << 15):
            break
        prec += 1
    v = wf * float(1 << prec)
    wi = np.where(v < 0, np.trunc(v - 0.5), np.trunc(v + 0.5)).astype(np.int16)
    return xmin, xsize, wi, prec


def _pass(img: np.ndarray, axis: int, out_size: int) -> np.ndarray:
    """One separable pass over `axis` of a uint8 [H,W,C] image."""
    in_size = img.shape[axis]
    if in_size == out_size:
        return img
    xmin, xsize, wi, prec = aa_weights_int16(in_size, out_size)
    src = np.moveaxis(img, axis, 0).astype(np.int32)
    out = np.empty((out_size,) + src.shape[1:], np.uint8)
    for i in range(out_size):
        n = int(xsize[i])
        acc = np.full(src.shape[1:], 1 << (prec - 1), np.int64)
        acc += np.tensordot(wi[i, :n].astype(np.int64), src[xmin[i]:xmin[i] + n], axes=(0, 0))
        out[i] = np.clip(acc >> prec, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def resize_bicubic_aa_u8(img_hwc: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """uint8 [H,W,C] -> uint8 [out_h,out_w,C]; horizontal pass, uint8 intermediate, vertical pass."""
    assert img_hwc.dtype == np.uint8 and img_hwc.ndim == 3
    return _pass(_pass(img_hwc, 1, out_w), 0, out_h)


def fused_mean_std(mean, std, rescale_factor: float = 1 / 255):
    """`_fuse_mean_std_and_rescale_factor` (image_processing_backends.py:303-306): float32 tensor * python float."""
    m = np.asarray(mean, np.float32) * np.float32(1.0 / rescale_factor)
    s = np.asarray(std, np.float32) * np.float32(1.0 / rescale_factor)
    return m.astype(np.float32), s.astype(np.float32)


def clip_preprocess(img_hwc: np.ndarray, size: int, mean=OPENAI_CLIP_MEAN, std=OPENAI_CLIP_STD,
                    square: bool = False) -> np.ndarray:
    """uint8 [H,W,3] -> float32 [3,size,size], the tensor `ImageProcessor.process` hands to the model.
    square=False: CLIP (shortest edge + center crop).  square=True: SigLIP (direct (size,size) resize, no crop)."""
    h, w = img_hwc.shape[:2]
    oh, ow = (size, size) if square else resize_output_size(h, w, size)
    r = resize_bicubic_aa_u8(img_hwc, oh, ow)
    top, left = center_crop_origin(oh, ow, size, size)
    r = r[top:top + size, left:left + size]
    m, s = fused_mean_std(mean, std)
    x = r.astype(np.float32).transpose(2, 0, 1)
    return ((x - m[:, None, None]) / s[:, None, None]).astype(np.float32)
