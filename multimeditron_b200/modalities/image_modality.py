"""B200-native image modality — drop-in for `ImageModality` (src/multimeditron/model/modalities/image_modality.py).

model_type "meditron_clip_b200": same config fields (hidden_size, clip_name, projection_type, use_2d_position_ids),
same processor contract, same state-dict keys (feature_extractor.vision_model.*, projector.projection.{0,2,4}.*),
same `forward(list[Tensor[3,h,w]]) -> Tensor[B, P, hidden]` — but the ViT forward, the projector and (optionally) the
splice run on the hand-written sm_100a kernels (multimeditron_b200.vit.TowerRunner).  Inference/frozen-embedder
path only in this round (the reference's ALIGNMENT/FULL backward is SURVEY.md §8(f)#2).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch

from ..configs import TOWERS, TowerConfig
from ..h2d import H2DUploader, HostStager
from ..preprocess import DevicePreprocessor, PackedImages
from ..vit import ProjectorWeights, TowerRunner, TowerWeights
from .base import AutoModality, BaseModality, BaseModalityConfig, BaseModalityProcessor
from .constants import MODALITY_VALUE_KEY, NUM_EMBEDDINGS_KEY, POSITION_IDS_KEY


def resolve_tower(clip_name) -> TowerConfig:
    """clip_name: a key of configs.TOWERS, a TowerConfig, a dict of its fields, or a local HF checkpoint dir."""
    if isinstance(clip_name, TowerConfig):
        return clip_name
    if isinstance(clip_name, dict):
        return TowerConfig(**clip_name)
    if clip_name in TOWERS:
        return TOWERS[clip_name]
    from transformers import AutoConfig  # local dir / hub name, like image_modality.py:76
    vc = AutoConfig.from_pretrained(clip_name).vision_config
    kind = "siglip" if "siglip" in type(vc).__name__.lower() else "clip"
    return TowerConfig(kind, vc.image_size, vc.patch_size, vc.hidden_size, vc.intermediate_size,
                       vc.num_hidden_layers, vc.num_attention_heads, vc.layer_norm_eps, vc.hidden_act)


class B200ImageConfig(BaseModalityConfig):
    def __init__(self, hidden_size: int = 4096, clip_name="openai/clip-vit-large-patch14",
                 projection_type: str = "mlp", use_2d_position_ids: bool = False,
                 device_preprocess: bool = False, **kwargs):
        super().__init__(modality_type="image", hidden_size=hidden_size, **kwargs)
        self.clip_name = clip_name
        self.projection_type = projection_type
        self.use_2d_position_ids = use_2d_position_ids
        # True: `process` hands over the raw uint8 [H,W,3] image and resize/crop/normalise run on the GPU
        # (mm_preprocess_images, bit-exact with the HF torchvision-backend processor) — SURVEY §8(f) rank 1
        self.device_preprocess = device_preprocess


class B200ImageProcessor(BaseModalityProcessor):
    """Same outputs as ImageProcessor.process (image_modality.py:83-110).  PIL images go through the HF image
    processor saved next to the checkpoint; tensors that are already [3,h,w] pixel values pass through."""

    def __init__(self, config):
        super().__init__(config)
        assert config.clip_name is not None, "clip_name must be specified in the config"
        self.tower = resolve_tower(config.clip_name)
        self._image_size = self.tower.grid
        self._num_patches_per_entry = self.tower.num_patches
        self.image_processor = None
        if isinstance(config.clip_name, str) and config.clip_name not in TOWERS:
            from transformers import AutoImageProcessor
            self.image_processor = AutoImageProcessor.from_pretrained(config.clip_name)

    def process(self, modality: Dict[str, Any]) -> Dict[str, Any]:
        out = modality.copy()
        image = modality[MODALITY_VALUE_KEY]
        if torch.is_tensor(image):
            out[MODALITY_VALUE_KEY] = image
        elif getattr(self.config, "device_preprocess", False):
            import numpy as np
            out[MODALITY_VALUE_KEY] = torch.from_numpy(np.array(image.convert("RGB"), dtype=np.uint8))  # convert_to_rgb
        else:
            if self.image_processor is None:
                raise ValueError("PIL input needs a checkpoint directory with an image processor (clip_name)")
            out[MODALITY_VALUE_KEY] = self.image_processor(images=image, return_tensors="pt")["pixel_values"][0]
        out[NUM_EMBEDDINGS_KEY] = self._num_patches_per_entry
        if self.config.use_2d_position_ids:
            g = self._image_size
            out[POSITION_IDS_KEY] = torch.stack(
                torch.meshgrid(torch.arange(g, dtype=torch.long), torch.arange(g, dtype=torch.long), indexing="ij"),
                dim=-1).reshape(self._num_patches_per_entry, 2)
        return out


@AutoModality.register("meditron_clip_b200")
class B200ImageModality(BaseModality):
    config_class = B200ImageConfig
    preprocessor_class = B200ImageProcessor

    def __init__(self, config: B200ImageConfig, state_dict: Optional[dict] = None, device=None):
        super().__init__(config)
        self.vision_tower_name = config.clip_name
        assert self.vision_tower_name is not None, "vision_tower_name must be specified in the config"
        self.tower = resolve_tower(config.clip_name)
        self.embedding_size = self.tower.hidden_size
        self._num_patches_per_entry = self.tower.num_patches
        self._runner: Optional[TowerRunner] = None
        self._stager: Optional[HostStager] = None
        self._pre: Optional[DevicePreprocessor] = None
        self._up: Optional[H2DUploader] = None
        self._frozen = True
        self.projector: Optional[torch.nn.Module] = None
        if state_dict is not None:
            self.load_reference_state_dict(state_dict, device or torch.device("cuda", torch.cuda.current_device()))

    # -- weights ------------------------------------------------------------------------------------------------
    def load_reference_state_dict(self, sd: dict, device) -> "B200ImageModality":
        """`sd` uses the reference's keys (ImageModality.state_dict())."""
        self._runner = TowerRunner(TowerWeights(self.tower, sd, device))
        return self

    @classmethod
    def from_reference(cls, ref_modality, device) -> "B200ImageModality":
        """Build from an instantiated reference `ImageModality` (same config, same parameters)."""
        c = ref_modality.config
        vc = ref_modality.feature_extractor.vision_model.config
        tower = TowerConfig("clip", vc.image_size, vc.patch_size, vc.hidden_size, vc.intermediate_size,
                            vc.num_hidden_layers, vc.num_attention_heads, vc.layer_norm_eps, vc.hidden_act)
        cfg = B200ImageConfig(hidden_size=c.hidden_size, clip_name=tower, projection_type=c.projection_type,
                              use_2d_position_ids=c.use_2d_position_ids)
        return cls(cfg, state_dict=ref_modality.state_dict(), device=device)

    @property
    def runner(self) -> TowerRunner:
        if self._runner is None:
            raise RuntimeError("B200ImageModality has no weights loaded (call load_reference_state_dict)")
        return self._runner

    # -- the plugin call ----------------------------------------------------------------------------------------
    def _preprocessor(self) -> DevicePreprocessor:
        """Device resize/crop/normalise.  A checkpoint directory carries its own `preprocessor_config.json`: the device
        path is built from THAT (mean/std, sizes, resample ...; unsupported settings raise).  Only the named synthetic
        towers of configs.TOWERS (no checkpoint, no processor file) use the architecture's stock defaults."""
        if self._pre is None:
            name = self.config.clip_name
            if isinstance(name, str) and name not in TOWERS:
                from transformers import AutoImageProcessor
                self._pre = DevicePreprocessor.from_hf_processor(AutoImageProcessor.from_pretrained(name), self.tower)
            else:
                self._pre = DevicePreprocessor.for_tower(self.tower)
            self._pre.uploader = self._uploader()
        return self._pre

    def _uploader(self) -> H2DUploader:
        if self._up is None:
            self._up = H2DUploader(self.runner.w.device)
        return self._up

    def _stack(self, inputs) -> torch.Tensor:
        """-> [B,3,S,S] pixels on the device.  Host inputs travel on the copy stream (h2d.H2DUploader); raw uint8
        images (list of [H,W,3], one [B,H,W,3] tensor, or PackedImages) are pre-processed on the device."""
        dev = self.runner.w.device
        if isinstance(inputs, PackedImages) or (torch.is_tensor(inputs) and inputs.dtype == torch.uint8):
            return self._preprocessor()(inputs, dev, torch.bfloat16)
        if torch.is_tensor(inputs):
            return self._uploader().upload(inputs)
        inputs = list(inputs)
        if len(inputs) and inputs[0].dtype == torch.uint8:   # raw [H,W,3] images: pre-process on the device
            return self._preprocessor()(inputs, dev, torch.bfloat16)
        if len(inputs) and inputs[0].is_cuda:
            return torch.stack(inputs, dim=0)
        if not inputs:
            return torch.empty(0, 3, self.tower.image_size, self.tower.image_size, device=dev)
        # the collator's format: a list of per-image CPU tensors -> pinned staging + one async H2D on the copy stream
        # instead of the reference's pageable, synchronous copy (:131-132)
        if self._stager is None:
            self._stager = HostStager(dev, self._uploader())
        return self._stager.stack_upload(inputs)

    def _release(self):
        if self._up is not None:
            self._up.release()

    @property
    def projector_trainable(self) -> bool:
        return self.projector is not None and any(p.requires_grad for p in self.projector.parameters())

    def forward(self, inputs: List[torch.Tensor]) -> torch.Tensor:
        """image_modality.py:130-137.  Frozen / no-grad: kernels only.  Projector unfrozen and autograd on: the tower
        runs on the kernels without a graph, the projector as autograd_ops.ProjectorFunction (forward + backward)."""
        if torch.is_grad_enabled() and self.projector_trainable:
            with torch.no_grad():
                px = self._stack(inputs)
                feats = self.runner.features(px).clone()
                self._release()
            from ..autograd_ops import projector_forward
            return projector_forward(feats, self.projector).view(px.shape[0], self.tower.num_patches, -1)
        with torch.no_grad():
            y = self.runner.forward(self._stack(inputs))
            self._release()
        return y

    @torch.no_grad()
    def forward_into(self, inputs, embedded_tokens: torch.Tensor, dest_rows: torch.Tensor) -> torch.Tensor:
        """forward + splice in one go: projected row k lands in embedded_tokens.view(-1,H)[dest_rows[k]]."""
        r = self.runner
        y = r.project(r.features(self._stack(inputs)), out=embedded_tokens, out_row_map=dest_rows)
        self._release()
        return y

    # -- training-mode switches (base.py:120-143) ---------------------------------------------------------------
    def freeze_modality_embedder(self):
        self._frozen = True

    def unfreeze_modality_embedder(self):
        raise NotImplementedError("ViT backward on the B200 path is not built (SURVEY.md §8(f)#2 lists it as optional); "
                                  "the reference's ALIGNMENT stage (projector only) is supported: unfreeze_projection()")

    def unfreeze_projection(self):
        """model.py:310-326 (freeze_for_alignment): expose the projector as nn.Parameters
        (`projector.projection.{0,2,4}.{weight,bias}` — the reference's state-dict keys) that SHARE storage with the
        kernel-side weights, so an optimizer's in-place updates are what the next forward multiplies by."""
        if self.projector is None:
            self.projector = _ProjectorParams(self.runner.w.proj)
        for p in self.projector.parameters():
            p.requires_grad = True

    def freeze_projection(self):
        if self.projector is not None:
            for p in self.projector.parameters():
                p.requires_grad = False


class _ProjectorParams(torch.nn.Module):
    """nn.Parameter view of ProjectorWeights (bf16, same storage), laid out like MLPProjector (projectors/mlp.py:33-39)."""

    def __init__(self, pw: ProjectorWeights):
        super().__init__()
        layers = []
        for w, b in ((pw.w0, pw.b0), (pw.w2, pw.b2), (pw.w4, pw.b4)):
            lin = torch.nn.Linear(w.shape[1], w.shape[0], device="meta")
            lin.weight = torch.nn.Parameter(w, requires_grad=False)
            lin.bias = torch.nn.Parameter(b, requires_grad=False)
            layers += [lin, torch.nn.GELU()]
        self.projection = torch.nn.Sequential(*layers[:-1])
