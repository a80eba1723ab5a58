"""Model-level entry of the hot path: `embed_modalities_with_text` on the B200 kernels.

Mirror of the part of `MultiModalModelForCausalLM` that is on the path (src/multimeditron/model/model.py):
  * processors()                      :379-380
  * _get_modality_by_name             :382-396   (KeyError / TypeError behaviour)
  * embed_modalities_with_text        :410-446   (embedding lookup + modality forward + splice)
The LLM itself (HF AutoModelForCausalLM) is out of scope: only its embedding table is touched by the path, so the
embedder takes that table.  `forward(...)` returns inputs_embeds, which the caller passes to its LLM as
`inputs_embeds=` exactly like model.py:517-526 does.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch

from . import ops
from .modalities.base import BaseModality, BaseModalityProcessor


class B200MultiModalEmbedder(torch.nn.Module):
    def __init__(self, embed_table: torch.Tensor, modalities: List[BaseModality], attachment_token_idx: int,
                 processors: Optional[Dict[str, BaseModalityProcessor]] = None, fuse_splice: bool = True):
        super().__init__()
        self.embed_table = embed_table  # [V, H] on the GPU, LLM dtype
        self.attachment_token_idx = int(attachment_token_idx)
        self.fuse_splice = fuse_splice
        self.modalities_by_type: Dict[str, BaseModality] = {}
        self.processors_by_type: Dict[str, BaseModalityProcessor] = dict(processors or {})
        for m in modalities:
            t = m.config.modality_type
            if t in self.modalities_by_type:  # model.py:275-278
                raise ValueError(f"Modality type {t} has already been registered")
            self.modalities_by_type[t] = m
        self.modalities_with_projection = torch.nn.ModuleList(modalities)
        self._status = None

    # model.py:379-380
    def processors(self) -> Dict[str, BaseModalityProcessor]:
        return self.processors_by_type

    # model.py:382-396
    def _get_modality_by_name(self, name: str) -> BaseModality:
        if name not in self.modalities_by_type:
            raise KeyError(f"Modality {name} not found in registered modalities")
        m = self.modalities_by_type[name]
        if not isinstance(m, BaseModality):
            raise TypeError(f"Registered modality {name} is not a BaseModality")
        return m

    def _host_stager(self, dev):
        from .h2d import HostTensorStager
        if getattr(self, "_stager", None) is None or self._stager.device != torch.device(dev):
            self._stager = HostTensorStager(dev)
        return self._stager

    def status(self) -> torch.Tensor:
        if self._status is None or self._status.device != self.embed_table.device:
            self._status = ops.new_status(self.embed_table.device)
        return self._status

    def check(self):
        """Lazy (one sync) check of data-dependent errors; raises IndexError / RuntimeError like the reference.
        The status word is cleared whether or not the check raises, so one bad batch cannot poison later checks.
        NOTE: for DEVICE-resident input_ids the result of `embed_modalities_with_text` is only valid once this
        check has passed (rows may be missing after a count mismatch); host-resident ids are verified eagerly."""
        if self._status is not None:
            try:
                ops.raise_for_status(self._status)
            finally:
                self._status.zero_()

    def _eager_host_checks(self, input_ids: torch.Tensor, stacks, positions_from: str):
        """The collator hands over HOST ids: verify what the reference would trip over BEFORE launching anything
        (same exception classes as nn.Embedding / index_put_, model.py:433,441) — no device round trip needed."""
        V, H = self.embed_table.shape
        if input_ids.numel() and (int(input_ids.min()) < 0 or int(input_ids.max()) >= V):
            raise IndexError("index out of range in self")
        if positions_from == "device":
            have = int((input_ids == self.attachment_token_idx).sum())
            want = 0
            for name, stack in stacks.items():
                m = self._get_modality_by_name(name)
                want += len(stack) * getattr(m, "rows_per_entry", m._num_patches_per_entry)
            if have != want:
                raise RuntimeError(f"shape mismatch: value tensor of shape [{want}, {H}] cannot be broadcast to "
                                   f"indexing result of shape [{have}, {H}]")

    def embed_modalities_with_text(self, input_ids: torch.Tensor, processed_multimodal_inputs: Dict[str, Any],
                                   positions_from: str = "collator", eager_checks: bool = True,
                                   skip_attachment_rows: Optional[bool] = None) -> torch.Tensor:
        """model.py:410-446.  positions_from:
          'collator' — use processed_multimodal_inputs['batch_idx'/'token_range'] as the reference does;
          'device'   — ignore them and derive the positions from input_ids with the device prefix sum
                       (valid for a single modality type); no index H2D at all.
        skip_attachment_rows (device positions only): do not copy the attachment token's own embedding row into the
          positions the splice overwrites anyway.  None (default) = only when the eager host check has just proved
          that the attachment-token count matches the modality rows (host ids, eager_checks=True); with a lazy check
          (device-resident ids or eager_checks=False) the rows ARE copied, so that on a count mismatch the result holds
          real embedding rows like the reference's (model.py:433) rather than uninitialised memory until check()
          raises.  True = the caller vouches for the counts and accepts that contract (bench.py's device-resident leg).
        Autograd: runs without a graph unless a modality's projector is trainable (`unfreeze_projection()`), in which
        case the projector and the splice are differentiable (autograd_ops); the embedding table is treated as frozen
        and a table that requires grad raises instead of losing its gradient silently."""
        dev = self.embed_table.device
        stacks = processed_multimodal_inputs["stacked"]
        if positions_from == "device" and len(stacks) > 1:
            raise ValueError("device-side positions need a single modality type")
        if torch.is_grad_enabled() and self.embed_table.requires_grad:
            raise NotImplementedError("the embedding table requires grad: the B200 splice path does not produce that "
                                      "gradient (freeze the LLM, or run under torch.no_grad())")
        host_checked = eager_checks and not input_ids.is_cuda
        if host_checked:
            self._eager_host_checks(input_ids, stacks, positions_from)
        if skip_attachment_rows is None:
            skip_attachment_rows = host_checked
        ids = self._host_stager(dev).upload(input_ids)      # pageable host ids -> pinned staging -> async H2D
        st = self.status()
        B, S = ids.shape
        H = self.embed_table.shape[1]
        trains = {n: torch.is_grad_enabled() and getattr(self._get_modality_by_name(n), "projector_trainable", False)
                  for n in stacks}
        with torch.no_grad():
            # device positions cover exactly the ids == attachment_id rows: those need no embedding copy (:433)
            skip = self.attachment_token_idx if (positions_from == "device" and stacks and skip_attachment_rows) else None
            embedded = ops.embed_gather(self.embed_table.detach(), ids, status=st, skip_id=skip)   # :433
        for name, stack in stacks.items():
            modality = self._get_modality_by_name(name)
            with torch.no_grad():
                if positions_from == "device":
                    n_rows = len(stack) * getattr(modality, "rows_per_entry", modality._num_patches_per_entry)
                    dest_rows, _, _ = ops.attach_positions(ids, self.attachment_token_idx, expected_rows=n_rows,
                                                           want_src_of_pos=False, status=st)
                else:
                    bi = self._host_stager(dev).upload(processed_multimodal_inputs["batch_idx"][name])
                    tr = self._host_stager(dev).upload(processed_multimodal_inputs["token_range"][name])
                    dest_rows = None
            fused = (self.fuse_splice and hasattr(modality, "forward_into") and embedded.dtype == torch.bfloat16
                     and not trains[name])
            if fused and dest_rows is not None:
                modality.forward_into(stack, embedded, dest_rows)                         # :439 + :441-444 fused
                continue
            proj = modality(stack)                                                        # :439
            proj2d = proj.reshape(-1, proj.shape[-1]).to(embedded.dtype)
            if proj.requires_grad:
                from .autograd_ops import splice_rows
                rows = dest_rows if dest_rows is not None else (bi * S + tr).to(torch.int32)
                if proj2d.shape[0] != rows.numel():
                    raise RuntimeError("shape mismatch: number of modality embeddings != attachment positions")
                embedded = splice_rows(embedded.view(B * S, H), proj2d, rows, st).view(B, S, H)
            elif dest_rows is not None:
                if proj2d.shape[0] != dest_rows.numel():
                    raise RuntimeError("shape mismatch: number of modality embeddings != attachment positions")
                with torch.no_grad():
                    ops.splice_scatter(embedded.view(B * S, H), proj2d.contiguous(), dest_rows, status=st)
            else:
                with torch.no_grad():
                    ops.splice_scatter_idx(embedded, proj2d.contiguous(), bi, tr, status=st)  # :441-444
        return embedded

    @torch.no_grad()
    def position_ids(self, input_ids: torch.Tensor, attention_mask: torch.Tensor, use_2d: bool = False,
                     modality: str = "image") -> torch.Tensor:
        """The collator's `position_ids` (data_loader.py:155-188) built on the device from the ids and the mask:
        [B,S] int64, or [B,S,2] with the image grid ids when `use_2d` (tokens per image / grid from the modality)."""
        dev = self.embed_table.device
        ids, mask = input_ids.to(dev, non_blocking=True), attention_mask.to(dev, non_blocking=True)
        if not use_2d:
            return ops.position_ids(mask)
        m = self._get_modality_by_name(modality)
        P = m._num_patches_per_entry
        g = int(round(P ** 0.5))
        return ops.position_ids(mask, ids, self.attachment_token_idx, P, g, use_2d=True, status=self.status())

    def forward(self, input_ids: torch.Tensor, processed_multimodal_inputs: Dict[str, Any], **kw) -> torch.Tensor:
        return self.embed_modalities_with_text(input_ids, processed_multimodal_inputs, **kw)
