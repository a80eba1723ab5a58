"""Reference-side binding: plug the B200 path into an installed MultiMeditron (the drop-in switch of INTEGRATION.md).

`register_with_reference()` registers `model_type="meditron_clip_b200"` with the REFERENCE's own `AutoModality`
(src/multimeditron/model/modalities/base.py:179-196 — which also wires HF AutoConfig/AutoModel/AutoProcessor), as a
subclass of the reference `ImageModality` (image_modality.py:113-149): the parameter tree, state-dict keys, processor
and `from_pretrained`/`save_pretrained` behaviour are the reference's, only `forward` is replaced by the sm_100a
kernels.  `patch_embed_modalities_with_text(model)` swaps `model.py:410-446` for the device-prefix-sum + fused-splice
version.  Nothing here is imported by the product path unless the reference is installed.
"""
from __future__ import annotations

import types
from typing import Any, Dict

import torch

from . import ops
from .configs import TowerConfig
from .vit import TowerRunner, TowerWeights

MODEL_TYPE = "meditron_clip_b200"
_registered = {}


def _tower_from_hf(vc) -> TowerConfig:
    kind = "siglip" if "siglip" in type(vc).__name__.lower() else "clip"
    return TowerConfig(kind, vc.image_size, vc.patch_size, vc.hidden_size, vc.intermediate_size,
                       vc.num_hidden_layers, vc.num_attention_heads, vc.layer_norm_eps, vc.hidden_act)


def register_with_reference():
    """-> the registered class.  Idempotent.  Needs `multimeditron` (the reference) importable."""
    if MODEL_TYPE in _registered:
        return _registered[MODEL_TYPE]
    from multimeditron.model.modalities import AutoModality
    from multimeditron.model.modalities.image_modality import ImageConfig, ImageModality, ImageProcessor

    class B200RefImageConfig(ImageConfig):
        # explicit __init__: transformers 5.x re-wraps configs that do not define one and drops the fields that
        # ImageConfig sets before calling PretrainedConfig.__init__ (modality_type, hidden_size)
        def __init__(self, hidden_size: int = 4096, clip_name: str = "openai/clip-vit-large-patch14",
                     projection_type: str = "mlp", use_2d_position_ids: bool = False, **kwargs):
            super().__init__(hidden_size=hidden_size, clip_name=clip_name, projection_type=projection_type,
                             use_2d_position_ids=use_2d_position_ids, **kwargs)

    class B200RefImageModality(ImageModality):
        """Reference `ImageModality` whose forward runs on the B200 kernels (frozen / inference path)."""
        config_class = B200RefImageConfig
        preprocessor_class = ImageProcessor

        def _runner(self) -> TowerRunner:
            dev = self.feature_extractor.device
            # re-pack when the parameters were (re)loaded or moved: key on data_ptr/version of one big weight
            w = self.projector.projection[4].weight
            key = (str(dev), w.data_ptr(), w._version)
            if getattr(self, "_b200_key", None) != key:
                tower = _tower_from_hf(self.feature_extractor.vision_model.config)
                self._b200_runner = TowerRunner(TowerWeights(tower, self.state_dict(), dev))
                self._b200_key = key
            return self._b200_runner

        def _check_frozen(self):
            if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
                raise NotImplementedError(
                    "meditron_clip_b200 implements the forward (frozen embedder / inference) path; the backward of "
                    "the projector/ViT is SURVEY.md §8(f)#2.  Call under torch.no_grad() or freeze_all().")

        def forward(self, inputs) -> torch.Tensor:                       # image_modality.py:130-137
            self._check_frozen()
            r = self._runner()
            px = torch.stack(list(inputs), dim=0) if not torch.is_tensor(inputs) else inputs
            return r.forward(px.to(r.w.device, non_blocking=True))

        def forward_into(self, inputs, embedded_tokens: torch.Tensor, dest_rows: torch.Tensor) -> torch.Tensor:
            self._check_frozen()
            r = self._runner()
            px = torch.stack(list(inputs), dim=0) if not torch.is_tensor(inputs) else inputs
            return r.project(r.features(px.to(r.w.device, non_blocking=True)), out=embedded_tokens,
                             out_row_map=dest_rows)

    cls = AutoModality.register(MODEL_TYPE)(B200RefImageModality)
    _registered[MODEL_TYPE] = cls
    return cls


def patch_embed_modalities_with_text(model, positions_from: str = "device"):
    """Replace `model.embed_modalities_with_text` (model.py:410-446) of a reference MultiModalModelForCausalLM
    instance: embedding lookup by `mm_embed_gather`, positions by the device prefix sum (or the collator's vectors),
    splice fused into the projector's last GEMM epilogue when the modality offers `forward_into`."""
    att = None
    for name in ("attachment_token_idx",):
        att = getattr(model.config, name, None) if att is None else att
    state: Dict[str, Any] = {"status": None}

    def embed_modalities_with_text(self, input_ids, processed_multimodal_inputs, attachment_token_idx=att):
        table = self.model.get_input_embeddings().weight
        dev = table.device
        ids = input_ids.to(dev, non_blocking=True)
        if state["status"] is None or state["status"].device != dev:
            state["status"] = ops.new_status(dev)
        st = state["status"]
        B, S = ids.shape
        H = table.shape[1]
        stacks = processed_multimodal_inputs["stacked"]
        embedded = ops.embed_gather(table.detach(), ids, status=st)
        use_dev = positions_from == "device" and len(stacks) == 1 and attachment_token_idx is not None
        for modality_name, stack in stacks.items():
            modality = self._get_modality_by_name(modality_name)
            if use_dev:
                n_rows = len(stack) * modality._num_patches_per_entry
                dest_rows, _, _ = ops.attach_positions(ids, attachment_token_idx, expected_rows=n_rows,
                                                       want_src_of_pos=False, status=st)
                if hasattr(modality, "forward_into") and embedded.dtype == torch.bfloat16:
                    modality.forward_into(stack, embedded, dest_rows)
                    continue
                proj = modality(stack)
                ops.splice_scatter(embedded.view(B * S, H), proj.reshape(-1, H).contiguous(), dest_rows, status=st)
            else:
                proj = modality(stack)
                bi = processed_multimodal_inputs["batch_idx"][modality_name].to(dev, non_blocking=True)
                tr = processed_multimodal_inputs["token_range"][modality_name].to(dev, non_blocking=True)
                ops.splice_scatter_idx(embedded, proj.reshape(-1, H).contiguous(), bi, tr, status=st)
        return embedded

    model.embed_modalities_with_text = types.MethodType(embed_modalities_with_text, model)
    model.b200_check = lambda: ops.raise_for_status(state["status"]) if state["status"] is not None else None
    return model
