"""Device image pre-processing: raw uint8 HWC images -> the `[B,3,S,S]` pixel tensor of the image embedder.

Host side of `mm_preprocess_images` (include/mm_b200.h).  It restates only the *geometry* the reference's image
processor decides on the host — the resize target (`get_resize_output_image_size`, HF:image_transforms.py:295-299)
and the center-crop origin (HF:image_processing_backends.py:360-361) — and ships raw bytes: 4x fewer PCIe bytes than
the fp32 `pixel_values` the reference uploads (image_modality.py:131-132), and no PIL/torchvision work on the CPU.
All arithmetic (tap weights, the two integer passes, normalisation) runs in the sm_100a kernels.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch

from . import ops

OPENAI_CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)   # HF:image_utils.py OPENAI_CLIP_MEAN / _STD
OPENAI_CLIP_STD = (0.26862954, 0.26130258, 0.27577711)
DESC_FIELDS = 8  # src_off, H, W, out_h, out_w, crop_top, crop_left, tmp_off


def resize_target(h: int, w: int, size: int, square: bool) -> tuple[int, int]:
    """CLIP: shortest edge -> size, long edge int(size*long/short); SigLIP: (size, size)."""
    if square:
        return size, size
    short, long = (w, h) if w <= h else (h, w)
    new_long = int(size * long / short)
    return (new_long, size) if w <= h else (size, new_long)


def taps_needed(in_size: int, out_size: int) -> int:
    scale = in_size / out_size
    support = 2.0 * scale if scale >= 1.0 else 2.0
    return int(math.ceil(np.float32(support))) * 2 + 1


@dataclass
class Plan:
    desc: np.ndarray      # int64 [B, 8]
    src_bytes: int
    tmp_bytes: int
    kmax: int
    max_h: int
    max_w: int


def plan_batch(shapes, size: int, square: bool) -> Plan:
    """shapes: [(H, W)] of the raw images.  Pure host logic (tested on CPU)."""
    desc = np.zeros((len(shapes), DESC_FIELDS), np.int64)
    src = tmp = 0
    kmax, mh, mw = 5, 1, 1
    for i, (h, w) in enumerate(shapes):
        if h <= 0 or w <= 0:
            raise ValueError(f"image {i} has an empty shape {(h, w)}")
        oh, ow = resize_target(h, w, size, square)
        if oh < size or ow < size:
            raise ValueError(f"image {i}: resized shape {(oh, ow)} smaller than the crop {size}")
        top, left = int((oh - size) / 2.0), int((ow - size) / 2.0)
        desc[i] = (src, h, w, oh, ow, top, left, tmp)
        src += (h * w * 3 + 3) & ~3            # keep every image 4-byte aligned
        tmp += h * size * 3
        kmax = max(kmax, taps_needed(h, oh), taps_needed(w, ow))
        mh, mw = max(mh, h), max(mw, w)
    return Plan(desc, src, tmp, kmax, mh, mw)


class PackedImages:
    """A ragged batch of raw images packed once: `buf` (uint8, pinned host or device) = images at their descriptor
    offsets followed by the int64 descriptor table; what a data loader would hand over instead of fp32 pixel tensors."""

    def __init__(self, buf: torch.Tensor, plan: Plan, desc_off: int):
        self.buf, self.plan, self.desc_off = buf, plan, desc_off

    def __len__(self) -> int:
        return self.plan.desc.shape[0]

    def to(self, device) -> "PackedImages":
        return PackedImages(self.buf.to(device, non_blocking=True), self.plan, self.desc_off)

    @property
    def nbytes(self) -> int:
        return self.buf.numel()


class DevicePreprocessor:
    """`pre(images) -> Tensor[B,3,S,S]` on the device; images are uint8 [H,W,3] tensors / arrays (CPU or CUDA)."""

    def __init__(self, size: int, mean=OPENAI_CLIP_MEAN, std=OPENAI_CLIP_STD, square: bool = False,
                 rescale_factor: float = 1 / 255):
        self.size, self.square = int(size), bool(square)
        # `_fuse_mean_std_and_rescale_factor` (HF:image_processing_backends.py:303-306): fp32 tensor * python float
        k = 1.0 / rescale_factor
        self.mean255 = (torch.tensor(mean, dtype=torch.float32) * k).tolist()
        self.std255 = (torch.tensor(std, dtype=torch.float32) * k).tolist()
        self._staging = None
        self._event = None
        self._bplan = None
        self.uploader = None     # optional h2d.H2DUploader: host buffers then travel on its copy stream

    def _h2d(self, host: torch.Tensor, device) -> torch.Tensor:
        if self.uploader is not None:
            return self.uploader.upload(host)
        return host.to(device, non_blocking=True)

    @classmethod
    def for_tower(cls, tower) -> "DevicePreprocessor":
        if tower.kind == "siglip":   # SiglipImageProcessor: (S,S) resize, no crop, mean = std = 0.5
            return cls(tower.image_size, (0.5, 0.5, 0.5), (0.5, 0.5, 0.5), square=True)
        return cls(tower.image_size)

    @classmethod
    def from_hf_processor(cls, ip, tower=None) -> "DevicePreprocessor":
        """Build from the checkpoint's HF image processor (`preprocessor_config.json`), so that the device path
        reproduces exactly what `ImageProcessor.process` (image_modality.py:96) would have produced on the CPU for THIS
        checkpoint.  Every setting the kernels do not implement raises instead of silently falling back to the
        OpenAI-CLIP defaults: resample != bicubic, do_resize/do_rescale/do_normalize off, non-square crops, a crop size
        different from the resize target (CLIP) or any crop at all for the square-resize processors (SigLIP), padding."""
        def get(name, default=None):
            return getattr(ip, name, default)

        def side(d, what):
            if d is None:
                return None
            if isinstance(d, int):
                return ("edge", int(d))
            d = dict(d)
            if "shortest_edge" in d and len(d) == 1:
                return ("edge", int(d["shortest_edge"]))
            if "height" in d and "width" in d:
                if int(d["height"]) != int(d["width"]):
                    raise ValueError(f"device preprocessing needs a square {what}, got {d}")
                return ("square", int(d["height"]))
            raise ValueError(f"unsupported {what} specification {d}")

        for flag in ("do_resize", "do_rescale", "do_normalize"):
            if get(flag, True) is False:
                raise ValueError(f"device preprocessing does not implement {flag}=False (image processor {type(ip).__name__})")
        if get("do_pad", False):
            raise ValueError("device preprocessing does not implement do_pad=True")
        rs = get("resample", 3)
        rs = int(getattr(rs, "value", rs)) if rs is not None else 3
        if rs != 3:      # PIL.Image.Resampling.BICUBIC == 3
            raise ValueError(f"device preprocessing implements bicubic resampling only (resample={rs})")
        size = side(get("size"), "size")
        if size is None:
            raise ValueError("image processor has no `size`")
        crop = side(get("crop_size"), "crop_size") if get("do_center_crop", False) else None
        if size[0] == "edge":          # CLIP style: shortest edge -> S, then center crop S x S
            if crop is None or crop[1] != size[1]:
                raise ValueError(f"shortest-edge resize needs a center crop of the same size (size={size[1]}, crop={crop})")
            square, out = False, size[1]
        else:                          # SigLIP style: resize to (S, S), no crop
            if crop is not None and crop[1] != size[1]:
                raise ValueError(f"square resize followed by a different center crop is not implemented (size={size[1]}, crop={crop[1]})")
            square, out = True, size[1]
        if tower is not None and out != tower.image_size:
            raise ValueError(f"image processor produces {out}px images, the tower expects {tower.image_size}px")
        mean, std = get("image_mean"), get("image_std")
        if mean is None or std is None or len(mean) != 3 or len(std) != 3:
            raise ValueError("image processor needs 3-channel image_mean / image_std")
        return cls(out, tuple(float(m) for m in mean), tuple(float(v) for v in std), square=square,
                   rescale_factor=float(get("rescale_factor", 1 / 255)))

    def _batched(self, images: torch.Tensor, device, out_dtype) -> torch.Tensor:
        """One uint8 [B,H,W,3] tensor (same-size images, e.g. a pinned host batch): no packing copy — the tensor itself
        is the source buffer, only the small descriptor table is staged."""
        B, H, W, C = images.shape
        if images.dtype != torch.uint8 or C != 3:
            raise ValueError(f"raw images must be uint8 [B,H,W,3], got {images.dtype} {tuple(images.shape)}")
        if B == 0:
            return torch.empty(0, 3, self.size, self.size, dtype=out_dtype, device=device)
        if (H * W * 3) % 4:
            return self([images[i] for i in range(B)], device, out_dtype)   # keep every image 4-byte aligned
        key = (B, H, W)
        if self._bplan is None or self._bplan[0] != key:
            plan = plan_batch([(H, W)] * B, self.size, self.square)
            desc = torch.from_numpy(plan.desc).pin_memory()
            self._bplan = (key, plan, desc, desc.to(device))              # descriptors of a fixed shape never change
        _, plan, _, desc_dev = self._bplan
        src = self._h2d(images.contiguous(), device).view(-1)
        return ops.preprocess_images(src, desc_dev, self.size, plan.kmax, plan.max_h, plan.max_w, plan.tmp_bytes,
                                     self.mean255, self.std255, out_dtype)

    def pack(self, images, pin: bool = True) -> PackedImages:
        """Host-side packing of a ragged list of uint8 [H,W,3] images (done once per batch, e.g. by the collator)."""
        images = [torch.from_numpy(np.ascontiguousarray(im)) if isinstance(im, np.ndarray) else im for im in images]
        for im in images:
            if im.dtype != torch.uint8 or im.dim() != 3 or im.shape[2] != 3:
                raise ValueError(f"raw images must be uint8 [H,W,3], got {im.dtype} {tuple(im.shape)}")
        plan = plan_batch([(im.shape[0], im.shape[1]) for im in images], self.size, self.square)
        desc_off = (plan.src_bytes + 7) & ~7
        buf = torch.empty(desc_off + plan.desc.nbytes, dtype=torch.uint8, pin_memory=pin)
        for im, d in zip(images, plan.desc):
            buf[int(d[0]):int(d[0]) + im.numel()].copy_(im.contiguous().view(-1))
        buf[desc_off:].copy_(torch.from_numpy(plan.desc.view(np.uint8).reshape(-1)))
        return PackedImages(buf, plan, desc_off)

    def _packed(self, packed: PackedImages, device, out_dtype) -> torch.Tensor:
        if len(packed) == 0:
            return torch.empty(0, 3, self.size, self.size, dtype=out_dtype, device=device)
        plan = packed.plan
        buf = packed.buf if packed.buf.is_cuda else self._h2d(packed.buf, device)   # ONE H2D per batch
        desc = buf[packed.desc_off:].view(torch.int64).view(-1, DESC_FIELDS)
        return ops.preprocess_images(buf, desc, self.size, plan.kmax, plan.max_h, plan.max_w, plan.tmp_bytes,
                                     self.mean255, self.std255, out_dtype)

    def __call__(self, images, device, out_dtype=torch.float32) -> torch.Tensor:
        if isinstance(images, PackedImages):
            return self._packed(images, device, out_dtype)
        if torch.is_tensor(images) and images.dim() == 4:
            return self._batched(images, device, out_dtype)
        images = [torch.from_numpy(np.ascontiguousarray(im)) if isinstance(im, np.ndarray) else im for im in images]
        for im in images:
            if im.dtype != torch.uint8 or im.dim() != 3 or im.shape[2] != 3:
                raise ValueError(f"raw images must be uint8 [H,W,3], got {im.dtype} {tuple(im.shape)}")
        if not images:
            return torch.empty(0, 3, self.size, self.size, dtype=out_dtype, device=device)
        plan = plan_batch([(im.shape[0], im.shape[1]) for im in images], self.size, self.square)
        desc_off = (plan.src_bytes + 7) & ~7
        total = desc_off + plan.desc.nbytes
        # one pinned staging buffer holds the packed images AND the descriptor table: a single async H2D per batch
        if self._staging is None or self._staging.numel() < total:
            self._staging = torch.empty(max(total, 1 << 20), dtype=torch.uint8, pin_memory=True)
        if self._event is not None:
            self._event.synchronize()
        host = self._staging
        on_device = []
        for im, d in zip(images, plan.desc):
            n = im.numel()
            if im.is_cuda:
                on_device.append((im, int(d[0]), n))
            else:
                host[int(d[0]):int(d[0]) + n].copy_(im.contiguous().view(-1))
        host[desc_off:total].copy_(torch.from_numpy(plan.desc.view(np.uint8).reshape(-1)))
        buf = host[:total].to(device, non_blocking=True)
        self._event = torch.cuda.Event()
        self._event.record()
        for im, off, n in on_device:
            buf[off:off + n].copy_(im.contiguous().view(-1))
        desc = buf[desc_off:total].view(torch.int64).view(-1, DESC_FIELDS)
        return ops.preprocess_images(buf, desc, self.size, plan.kmax, plan.max_h, plan.max_w, plan.tmp_bytes,
                                     self.mean255, self.std255, out_dtype)

    def h2d_bytes(self, images) -> int:
        plan = plan_batch([(im.shape[0], im.shape[1]) for im in images], self.size, self.square)
        return ((plan.src_bytes + 7) & ~7) + plan.desc.nbytes
