"""B200-native mixture-of-experts image modalities — drop-ins for
  * `MOEImageModality`     (src/multimeditron/model/modalities/image_modality_moe.py:91-246,  "moe_meditron_clip")
  * `MOEImageModalityPEP`  (src/multimeditron/model/modalities/image_modality_moe_pep.py:91-288, "moe_meditron_clip_pep")

Reference behaviour reproduced:
  * EVERY expert runs on EVERY image; top-k is computed by the gating network and discarded;
  * `weighted_average`: shared-projector variant permutes the gating weights into expert order
    (image_modality_moe.py:171-177); the PEP variant does NOT (image_modality_moe_pep.py:212-216);
  * `sequence_append`: [B,E,P,*] flattened to [B,E*P,*] (:167-169 / :206-209);
  * state-dict keys `experts.<i>.*` (HF CLIPVisionTransformer), `projector.projection.*` (shared) or
    `projectors.<i>.projection.*` (PEP), `gating_network.resnet.*`.
The expert towers, projectors and all three fusions (`weighted_average`, `sequence_append`, `cross_attn` — the last
one inference-only: its backward is not written and `unfreeze_projection()` raises for it) run on the sm_100a kernels.
The gating network (ResNet-50, moe/gating.py:73-89) is handed in by the caller (`load_reference_state_dict(..., gating=)`
or `set_gating`): normally a `gating.B200ResNet50Gate` built from the `gating_network.resnet.*` entries of the same state
dict (row staging + tcgen05 GEMMs, BatchNorm/ReLU folded), but any callable `gating(pixels[B,3,h,w]) -> weights[B,E]`
(a torch module, precomputed weights) is accepted.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Sequence

import torch

from .. import ops
from ..configs import TOWERS, TowerConfig
from ..h2d import H2DUploader
from ..preprocess import DevicePreprocessor, PackedImages
from ..vit import FOLD_LN, GroupedTowerRunner, ProjectorWeights, TowerRunner, TowerWeights
from .base import AutoModality, BaseModality, BaseModalityConfig, BaseModalityProcessor
from .constants import MODALITY_VALUE_KEY, NUM_EMBEDDINGS_KEY
from .image_modality import resolve_tower


class B200MOEImageConfig(BaseModalityConfig):
    """Same fields as MOEImageConfig (image_modality_moe.py:11-52) / MOEImageConfigPEP."""

    def __init__(self, hidden_size: int = 1024, use_bias_proj: bool = True, expert_clip_names: Sequence = (),
                 image_processor="openai/clip-vit-large-patch14", gating_path: str = "", top_k_experts: int = 1,
                 projection_type: str = "mlp", generalist_idx: int = -1, fusion_method: str = "weighted_average",
                 cross_attn_heads: int = 8, gating_class_names: Optional[Sequence[str]] = None,
                 expert_towers: Optional[Dict[str, Any]] = None, **kwargs):
        super().__init__(modality_type="image", hidden_size=hidden_size, **kwargs)
        self.use_bias_proj = use_bias_proj
        self.expert_clip_names = list(expert_clip_names)
        self.top_k_experts = top_k_experts
        self.gating_path = gating_path
        self.projection_type = projection_type
        self.image_processor = image_processor
        self.generalist_idx = generalist_idx
        self.fusion_method = fusion_method
        self.cross_attn_heads = cross_attn_heads
        # GatingNetworkConfig.class_names (read at image_modality_moe.py:125); kept in the config here because the
        # gating network itself is supplied by the caller
        self.gating_class_names = list(gating_class_names) if gating_class_names else None
        # optional offline shape table: expert name -> configs.TOWERS key / TowerConfig fields (instead of resolving
        # the name through HF AutoConfig, which needs the checkpoint directory or the hub)
        self.expert_towers = dict(expert_towers) if expert_towers else None


class B200MOEImageConfigPEP(B200MOEImageConfig):
    pass


class B200MOEImageProcessor(BaseModalityProcessor):
    """MOEImageProcessor.process (image_modality_moe.py:55-87): num_embeddings = P * top_k_experts for
    `sequence_append`, P for `weighted_average` / `cross_attn`, ValueError otherwise."""

    def __init__(self, config):
        super().__init__(config)
        self.tower = resolve_tower(config.image_processor)
        self._num_patches_per_entry = self.tower.num_patches
        self.top_k_experts = config.top_k_experts
        self.fusion_method = config.fusion_method
        self.image_processor = None
        from ..configs import TOWERS
        if isinstance(config.image_processor, str) and config.image_processor not in TOWERS:
            from transformers import AutoImageProcessor
            self.image_processor = AutoImageProcessor.from_pretrained(config.image_processor)

    def process(self, modality: Dict[str, Any]) -> Dict[str, Any]:
        out = modality.copy()
        image = modality[MODALITY_VALUE_KEY]
        if not torch.is_tensor(image):
            if self.image_processor is None:
                raise ValueError("PIL input needs a checkpoint directory with an image processor")
            image = self.image_processor(images=image, return_tensors="pt")["pixel_values"][0]
        out[MODALITY_VALUE_KEY] = image
        if self.fusion_method == "sequence_append":
            out[NUM_EMBEDDINGS_KEY] = self._num_patches_per_entry * self.top_k_experts
        elif self.fusion_method in ("weighted_average", "cross_attn"):
            out[NUM_EMBEDDINGS_KEY] = self._num_patches_per_entry
        else:
            raise ValueError(f"Unknown fusion_method: {self.fusion_method}")
        return out


class _B200MOEBase(BaseModality):
    """Shared host logic of the two MoE modalities (expert loop, gating alignment, fusion dispatch)."""

    config_class = B200MOEImageConfig
    preprocessor_class = B200MOEImageProcessor
    per_expert_projector = False   # PEP subclass sets True

    def __init__(self, config, state_dict: Optional[dict] = None, device=None,
                 gating: Optional[Callable[[torch.Tensor], torch.Tensor]] = None):
        super().__init__(config)
        self.expert_names: List = list(config.expert_clip_names)
        assert len(self.expert_names) > 0, "config.expert_clip_names must be non-empty"
        table = getattr(config, "expert_towers", None) or {}
        self.towers: List[TowerConfig] = [resolve_tower(table.get(n, n) if isinstance(n, str) else n)
                                          for n in self.expert_names]
        t0 = self.towers[0]
        for t in self.towers[1:]:
            # image_modality_moe_pep.py:136-141 (and torch.stack at image_modality_moe.py:165 for the shared variant)
            assert t.patch_size == t0.patch_size and t.image_size == t0.image_size, \
                "sequence_append requires identical (image_size, patch_size) across experts."
            if not self.per_expert_projector:
                assert t.hidden_size == t0.hidden_size, "shared projector needs experts of one width"
        self._num_patches_per_entry = t0.num_patches
        self._native_embed_dim = t0.hidden_size
        self.generalist_idx = config.generalist_idx
        self.fusion_method = config.fusion_method
        if self.fusion_method.replace("-", "_") == "cross_attn":     # image_modality_moe.py:141 accepts both spellings
            self.fusion_method = "cross_attn"
        if self.fusion_method not in ("weighted_average", "sequence_append", "cross_attn"):
            raise ValueError(f"Unsupported fusion_method: {self.fusion_method}")
        # perm[class_idx] = expert_idx (image_modality_moe.py:125-138)
        names = getattr(config, "gating_class_names", None) or []
        if names:
            idx = {str(n): i for i, n in enumerate(map(str, self.expert_names))}
            try:
                perm = [idx[str(n)] for n in names]
            except KeyError as e:
                raise ValueError(f"Gating class name {e} not found in expert_clip_names: {self.expert_names}")
        else:
            perm = list(range(len(self.towers)))
        self._perm_list = perm
        self._gating = gating
        self._pre: Optional[DevicePreprocessor] = None
        self._runners: List[TowerRunner] = []
        self._projs: List[ProjectorWeights] = []
        self._maps: Dict[Any, torch.Tensor] = {}
        self._staging = None
        if state_dict is not None:
            self.load_reference_state_dict(state_dict, device or torch.device("cuda", torch.cuda.current_device()))

    @property
    def embedding_size(self) -> int:
        # shared: experts' native width (image_modality_moe.py:114-115); PEP: post-projection width (:251-253)
        return self.config.hidden_size if self.per_expert_projector else self._native_embed_dim

    @property
    def rows_per_entry(self) -> int:
        """Embedding rows one image contributes to the sequence (E*P for sequence_append)."""
        P = self._num_patches_per_entry
        return P * len(self.towers) if self.fusion_method == "sequence_append" else P

    # -- weights ------------------------------------------------------------------------------------------------
    def load_reference_state_dict(self, sd: dict, device, gating=None):
        dev = torch.device(device)
        pools: Dict[TowerConfig, dict] = {}
        self._runners = [TowerRunner(TowerWeights(t, sd, dev, vm_prefix=f"experts.{i}.", with_projector=False),
                                     shared_ws=pools.setdefault(t, {})) for i, t in enumerate(self.towers)]
        if self.per_expert_projector:
            self._projs = [ProjectorWeights(sd, dev, f"projectors.{i}.") for i in range(len(self.towers))]
        else:
            self._projs = [ProjectorWeights(sd, dev, "projector.")]
        self._device = dev
        # same-shaped experts: one grouped launch per layer op for all of them (vit.GroupedTowerRunner); used whenever
        # the batch allows it (B*T % 256 == 0), the per-expert runners otherwise.  MM_B200_MOE_GROUPED=0: A/B switch.
        import os
        self._grouped = None
        if (len(self.towers) > 1 and all(t == self.towers[0] for t in self.towers) and FOLD_LN
                and os.environ.get("MM_B200_MOE_GROUPED", "1") != "0"):
            self._grouped = GroupedTowerRunner([r.w for r in self._runners])
        self._perm = torch.tensor(self._perm_list, dtype=torch.int32, device=dev)
        self._xattn = None
        if self.fusion_method == "cross_attn":    # CrossAttention parameters (attention.py:36-42), k|v fused into one GEMM
            g = lambda n: sd["cross_attn." + n].detach().to(device=dev, dtype=torch.bfloat16).contiguous()  # noqa: E731
            self._xattn = dict(wq=g("q_proj.weight"), bq=g("q_proj.bias"),
                               wkv=torch.cat([g("k_proj.weight"), g("v_proj.weight")], 0).contiguous(),
                               bkv=torch.cat([g("k_proj.bias"), g("v_proj.bias")], 0).contiguous(),
                               wo=g("proj.weight"), bo=g("proj.bias"))
        if gating is not None:
            self._gating = gating
        return self

    def set_gating(self, gating: Callable[[torch.Tensor], torch.Tensor]):
        """gating(pixels [B,3,h,w] on the device) -> weights [B,E] (or the reference's (logits, topk, weights))."""
        self._gating = gating
        return self

    # -- helpers ------------------------------------------------------------------------------------------------
    def _stack(self, inputs) -> torch.Tensor:
        if not self._runners:
            raise RuntimeError("no weights loaded (call load_reference_state_dict)")
        raw_list = isinstance(inputs, (list, tuple)) and len(inputs) and inputs[0].dtype == torch.uint8
        if isinstance(inputs, PackedImages) or raw_list or (torch.is_tensor(inputs) and inputs.dtype == torch.uint8):
            # raw uint8 images (mixed resolutions): resize / crop / normalise on the device, bit-exact with the
            # experts' shared HF image processor (image_modality_moe.py:96-99 of the reference)
            if self._pre is None:
                ipn = getattr(self.config, "image_processor", None)
                if isinstance(ipn, str) and ipn not in TOWERS:    # a checkpoint directory: use ITS preprocessor_config
                    from transformers import AutoImageProcessor
                    self._pre = DevicePreprocessor.from_hf_processor(AutoImageProcessor.from_pretrained(ipn), self.towers[0])
                else:
                    self._pre = DevicePreprocessor.for_tower(self.towers[0])
                self._pre.uploader = H2DUploader(self._device)     # host bytes travel on a copy stream
            return self._pre(inputs, self._device, torch.bfloat16)
        if torch.is_tensor(inputs):
            return inputs.to(self._device, non_blocking=True)
        inputs = list(inputs)
        if inputs and inputs[0].is_cuda:
            return torch.stack(inputs, 0)
        return torch.stack(inputs, 0).pin_memory().to(self._device, non_blocking=True)

    def _weights(self, px: torch.Tensor) -> torch.Tensor:
        if self._gating is None:
            raise RuntimeError("no gating network set: call set_gating(callable) "
                               "(the ResNet-50 gate is not part of the B200 path, SURVEY.md §8(f)#3)")
        w = self._gating(px)
        if isinstance(w, (tuple, list)):
            w = w[2]
        E = len(self.towers)
        if w.shape != (px.shape[0], E):
            raise RuntimeError(f"gating returned weights of shape {tuple(w.shape)}, expected {(px.shape[0], E)}")
        return w.to(device=self._device, dtype=torch.float32).contiguous()

    def _append_map(self, B: int, e: int, E: Optional[int] = None) -> torch.Tensor:
        """compact row (b*P+p) of expert e -> row ((b*E+e)*P+p) of the [B,E,P,*] stack."""
        E = len(self.towers) if E is None else E
        key = (B, e, E)
        m = self._maps.get(key)
        if m is None:
            if len(self._maps) > 64:
                self._maps.clear()
            P = self._num_patches_per_entry
            r = torch.arange(B * P, device=self._device, dtype=torch.int64)
            m = (((r // P) * E + e) * P + r % P).to(torch.int32)
            self._maps[key] = m
        return m

    # -- the plugin call ------------------------------------------------------------------------------------------
    @property
    def projector_trainable(self) -> bool:
        mods = [m for m in (getattr(self, "projector", None), getattr(self, "projectors", None)) if m is not None]
        return any(p.requires_grad for m in mods for p in m.parameters())

    def forward(self, inputs) -> torch.Tensor:
        """-> [B, P, H] (weighted_average / cross_attn) or [B, E*P, H] (sequence_append), bf16.
        ALIGNMENT stage (`unfreeze_projection()`, autograd on): experts, gate and fusion run on the kernels without a
        graph, the projector(s) as `autograd_ops.ProjectorFunction` (forward + backward)."""
        if torch.is_grad_enabled() and self.projector_trainable:
            return self._forward_train(inputs)
        with torch.no_grad():
            return self._forward_infer(inputs)

    def _forward_infer(self, inputs) -> torch.Tensor:
        px = self._stack(inputs)
        B, E, P, H = px.shape[0], len(self.towers), self._num_patches_per_entry, self.config.hidden_size
        n_out = E * P if self.fusion_method == "sequence_append" else P
        out = torch.empty(B, n_out, H, dtype=torch.bfloat16, device=self._device)
        self._run(px, out.view(-1, H), None)
        self._release()
        return out

    @torch.no_grad()
    def forward_into(self, inputs, embedded_tokens: torch.Tensor, dest_rows: torch.Tensor) -> torch.Tensor:
        """forward + splice: output row k lands in embedded_tokens.view(-1,H)[dest_rows[k]] (model.py:441-444)."""
        px = self._stack(inputs)
        self._run(px, embedded_tokens.view(-1, embedded_tokens.shape[-1]), dest_rows)
        self._release()
        return embedded_tokens

    def _release(self):
        if self._pre is not None and self._pre.uploader is not None:
            self._pre.uploader.release()

    def _cross_attn_fuse(self, stack: torch.Tensor, w: torch.Tensor, B: int, P: int) -> torch.Tensor:
        """image_modality_moe.py:178-205 / image_modality_moe_pep.py:218-245 on `stack` [E, B*P, C]:
        generalist tokens (python index `generalist_idx`, -1 = last expert) query the gate-weighted tokens of every
        expert i with `i != generalist_idx` — the reference compares the RAW index, so with the default -1 the generalist
        is also one of the "specialists"; reproduced as is.  Gate weights: permuted to expert order (both variants do it
        here, :190 / :230), re-softmaxed over the specialists in the tensor dtype, multiplied into the tokens (bf16
        products), k|v projected by ONE GEMM straight into the [B, S*P, 2C] layout, attention (mm_cross_attention),
        output projection."""
        E, _, C = stack.shape
        gi = self.generalist_idx
        spec = [i for i in range(E) if i != gi]
        xa = self._xattn
        w_all = w.to(torch.bfloat16).index_select(-1, self._perm.long())               # [B, E] in expert order
        w_spec = torch.softmax(w_all[:, spec], dim=-1).float().contiguous()               # bf16 values, held as fp32
        S = len(spec)
        kv = torch.empty(B, S * P, 2 * C, dtype=torch.bfloat16, device=self._device)
        for j, e in enumerate(spec):
            ctx = ops.weighted_expert_sum(stack[e:e + 1], w_spec[:, j:j + 1].contiguous(), None, B, P)   # ctx * wj
            ops.gemm_bias_act(ctx, xa["wkv"], xa["bkv"], out=kv.view(-1, 2 * C), out_row_map=self._append_map(B, j, S))
        q = ops.gemm_bias_act(stack[gi], xa["wq"], xa["bq"])                              # [B*P, C]
        a = ops.cross_attention(q.view(B, P, C), kv, self.config.cross_attn_heads)
        return ops.gemm_bias_act(a.view(B * P, C), xa["wo"], xa["bo"])

    def _experts_and_gate(self, px: torch.Tensor, stack: torch.Tensor) -> torch.Tensor:
        """stack[e] = expert e's CLS-dropped features [B*P, D] (image_modality_moe.py:160-162); returns the gate's
        weights [B,E] (:156).  Same-shaped experts and an eligible batch: ONE grouped launch per layer op on the current
        stream, the gate on a side stream next to it; otherwise one stream per expert and the gate on the current one."""
        if self._grouped is not None and self._grouped.eligible(px.shape[0]):
            import os
            if os.environ.get("MM_B200_MOE_STREAMS", "1") == "0":      # A/B and per-op profiling: everything on one stream
                self._grouped.features(px, stack)
                self._joins = None
                return self._weights(px)
            cur = torch.cuda.current_stream(self._device)
            if getattr(self, "_gate_stream", None) is None:
                self._gate_stream = torch.cuda.Stream(device=self._device)
            fork = torch.cuda.Event()
            fork.record(cur)
            self._gate_stream.wait_event(fork)
            with torch.cuda.stream(self._gate_stream):
                w = self._weights(px)
                done = torch.cuda.Event()
                done.record(self._gate_stream)
            self._grouped.features(px, stack)
            cur.wait_event(done)
            # w was allocated from the gate stream's pool but is consumed on `cur`: tell the caching allocator, so the
            # block cannot be handed out again on the gate stream before cur's readers have run (today the next use of
            # the gate stream waits on a fork event recorded after them, but callers need not keep to that pattern)
            w.record_stream(cur)
            self._joins = None
            return w
        self._each_expert(lambda e, r: r.features(px, out=stack[e]))
        w = self._weights(px)
        self._join_experts()
        return w

    def _each_expert(self, fn):
        """Run fn(e, runner) for every expert, expert e on its own CUDA stream (fork from / join into the current
        stream).  Each expert's GEMMs are persistent kernels with a short tail wave (cfg5: 150 tiles on 74 CTA pairs =
        2.03 waves); side by side, expert e+1's CTAs start on the SMs expert e's tail leaves idle, and the gating network
        (current stream) overlaps the experts.  MM_B200_MOE_STREAMS=0 serialises them on the current stream (A/B)."""
        import os
        if os.environ.get("MM_B200_MOE_STREAMS", "1") == "0" or len(self._runners) == 1:
            for e, r in enumerate(self._runners):
                fn(e, r)
            return
        cur = torch.cuda.current_stream(self._device)
        if getattr(self, "_streams", None) is None or len(self._streams) != len(self._runners):
            self._streams = [torch.cuda.Stream(device=self._device) for _ in self._runners]
        fork = torch.cuda.Event()
        fork.record(cur)
        joins = []
        for e, r in enumerate(self._runners):
            st = self._streams[e]
            st.wait_event(fork)
            with torch.cuda.stream(st):
                fn(e, r)
                ev = torch.cuda.Event()
                ev.record(st)
                joins.append(ev)
        self._joins = joins

    def _join_experts(self):
        cur = torch.cuda.current_stream(self._device)
        for ev in getattr(self, "_joins", None) or []:
            cur.wait_event(ev)
        self._joins = None

    def _run(self, px: torch.Tensor, out2d: torch.Tensor, dest_rows: Optional[torch.Tensor]):
        raise NotImplementedError

    def _forward_train(self, inputs) -> torch.Tensor:
        raise NotImplementedError

    # -- training-mode switches (image_modality_moe.py:222-246, image_modality_moe_pep.py:263-288) ------------------
    def freeze_modality_embedder(self):
        self.modality_frozen = True

    def unfreeze_modality_embedder(self):
        raise NotImplementedError("ViT / gating backward is not part of the B200 path (SURVEY.md §8(f)#2 lists it as "
                                  "optional); the ALIGNMENT stage (projectors only) is: unfreeze_projection()")

    def unfreeze_projection(self):
        """model.py:310-326 (freeze_for_alignment): expose the projector(s) as nn.Parameters that SHARE storage with the
        kernel-side weights (`projector.projection.*` for the shared variant, `projectors.<e>.projection.*` for PEP — the
        reference's state-dict keys), so an optimizer's in-place updates are what the next forward multiplies by."""
        from .image_modality import _ProjectorParams
        if self.fusion_method == "cross_attn":
            raise NotImplementedError("cross_attn fusion keeps its CrossAttention parameters trainable in the reference's "
                                      "ALIGNMENT stage (they are never frozen): its backward is not built")
        if self.per_expert_projector:
            if getattr(self, "projectors", None) is None:
                self.projectors = torch.nn.ModuleList([_ProjectorParams(pw) for pw in self._projs])
            mods = self.projectors
        else:
            if getattr(self, "projector", None) is None:
                self.projector = _ProjectorParams(self._projs[0])
            mods = self.projector
        for p in mods.parameters():
            p.requires_grad = True

    def freeze_projection(self):
        for m in (getattr(self, "projector", None), getattr(self, "projectors", None)):
            if m is not None:
                for p in m.parameters():
                    p.requires_grad = False


@AutoModality.register("moe_meditron_clip_b200")
class B200MOEImageModality(_B200MOEBase):
    """Shared projector, fusion in the experts' feature space (image_modality_moe.py:152-210)."""
    config_class = B200MOEImageConfig
    preprocessor_class = B200MOEImageProcessor
    per_expert_projector = False

    def _run(self, px, out2d, dest_rows):
        B, E, P = px.shape[0], len(self.towers), self._num_patches_per_entry
        D = self._native_embed_dim
        r0, proj = self._runners[0], self._projs[0]
        if self.fusion_method == "weighted_average":
            stack = torch.empty(E, B * P, D, dtype=torch.bfloat16, device=self._device)
            w = self._experts_and_gate(px, stack)                                  # :156, :160-162 (every expert)
            fused = ops.weighted_expert_sum(stack, w, self._perm, B, P)            # :171-177 (permuted weights)
        elif self.fusion_method == "cross_attn":                                   # :178-205
            stack = torch.empty(E, B * P, D, dtype=torch.bfloat16, device=self._device)
            w = self._experts_and_gate(px, stack)
            fused = self._cross_attn_fuse(stack, w, B, P)
        else:                                                                      # sequence_append :167-169
            fused = torch.empty(B * E * P, D, dtype=torch.bfloat16, device=self._device)
            maps = [self._append_map(B, e) for e in range(E)]
            self._each_expert(lambda e, r: r.features(px, out=fused, out_row_map=maps[e]))
            self._join_experts()
        r0.project(fused, proj, out=out2d, out_row_map=dest_rows)                  # :207


    def _forward_train(self, inputs) -> torch.Tensor:
        """Shared projector after the fusion: everything before it runs without a graph."""
        from ..autograd_ops import projector_forward
        with torch.no_grad():
            px = self._stack(inputs)
            B, E, P = px.shape[0], len(self.towers), self._num_patches_per_entry
            D = self._native_embed_dim
            if self.fusion_method == "weighted_average":
                stack = torch.empty(E, B * P, D, dtype=torch.bfloat16, device=self._device)
                w = self._experts_and_gate(px, stack)
                fused = ops.weighted_expert_sum(stack, w, self._perm, B, P)
            else:
                fused = torch.empty(B * E * P, D, dtype=torch.bfloat16, device=self._device)
                maps = [self._append_map(B, e) for e in range(E)]
                self._each_expert(lambda e, r: r.features(px, out=fused, out_row_map=maps[e]))
                self._join_experts()
            self._release()
        return projector_forward(fused, self.projector).view(B, -1, self.config.hidden_size)


@AutoModality.register("moe_meditron_clip_pep_b200")
class B200MOEImageModalityPEP(_B200MOEBase):
    """One projector per expert, fusion in the LLM's space (image_modality_moe_pep.py:191-249)."""
    config_class = B200MOEImageConfigPEP
    preprocessor_class = B200MOEImageProcessor
    per_expert_projector = True

    def _run(self, px, out2d, dest_rows):
        B, E, P, H = px.shape[0], len(self.towers), self._num_patches_per_entry, self.config.hidden_size
        if self.fusion_method == "weighted_average":
            stack = torch.empty(E, B * P, H, dtype=torch.bfloat16, device=self._device)
            self._each_expert(lambda e, r: r.project(r.features(px), self._projs[e], out=stack[e]))   # :198-203
            w = self._weights(px)                                                  # :194 (overlaps the experts)
            self._join_experts()
            fused = ops.weighted_expert_sum(stack, w, None, B, P)                  # :212-216 (NO permutation)
            if dest_rows is None:
                out2d.copy_(fused)
            else:
                ops.splice_scatter(out2d, fused, dest_rows)
        elif self.fusion_method == "cross_attn":                                   # :218-245 (in the projected space)
            stack = torch.empty(E, B * P, H, dtype=torch.bfloat16, device=self._device)
            self._each_expert(lambda e, r: r.project(r.features(px), self._projs[e], out=stack[e]))
            w = self._weights(px)
            self._join_experts()
            fused = self._cross_attn_fuse(stack, w, B, P)
            if dest_rows is None:
                out2d.copy_(fused)
            else:
                ops.splice_scatter(out2d, fused, dest_rows)
        else:                                                                      # sequence_append :206-209
            maps = []
            for e in range(E):
                m = self._append_map(B, e)
                maps.append(m if dest_rows is None else dest_rows[m.long()].contiguous())
            self._each_expert(lambda e, r: r.project(r.features(px), self._projs[e], out=out2d, out_row_map=maps[e]))
            self._join_experts()

    def _forward_train(self, inputs) -> torch.Tensor:
        """One trainable projector per expert, fusion AFTER them: features without a graph, projectors and the weighted sum
        (or the concatenation) under autograd."""
        from ..autograd_ops import projector_forward, weighted_expert_sum
        with torch.no_grad():
            px = self._stack(inputs)
            B, E, P, H = px.shape[0], len(self.towers), self._num_patches_per_entry, self.config.hidden_size
            feats = torch.empty(E, B * P, self._native_embed_dim, dtype=torch.bfloat16, device=self._device)
            if self.fusion_method == "weighted_average":
                w = self._experts_and_gate(px, feats)
            else:
                self._each_expert(lambda e, r: r.features(px, out=feats[e]))
                self._join_experts()
            self._release()
        ys = [projector_forward(feats[e], self.projectors[e]) for e in range(E)]             # [B*P, H] each
        if self.fusion_method == "weighted_average":
            return weighted_expert_sum(torch.stack(ys, 0), w, None, B, P).view(B, P, H)      # :212-216 (no permutation)
        return torch.stack([y.view(B, P, H) for y in ys], 1).reshape(B, E * P, H)            # :206-209

