"""Reference-side binding: plug the B200 path into an installed MultiMeditron (the drop-in switch of INTEGRATION.md).

`register_with_reference()` registers `model_type="meditron_clip_b200"` with the REFERENCE's own `AutoModality`
(src/multimeditron/model/modalities/base.py:179-196 — which also wires HF AutoConfig/AutoModel/AutoProcessor), as a
subclass of the reference `ImageModality` (image_modality.py:113-149): the parameter tree, state-dict keys, processor
and `from_pretrained`/`save_pretrained` behaviour are the reference's, only `forward` is replaced by the sm_100a
kernels.  `patch_embed_modalities_with_text(model, attachment_token_idx=...)` swaps `model.py:410-446` for the
device-prefix-sum + fused-splice version and wraps `generate` (model.py:528-640) in `torch.no_grad()`.
Nothing here is imported by the product path unless the reference is installed.

Gradient contract (the reference's freeze modes, model.py:310-377):
  * everything frozen / `torch.no_grad()` (inference, `generate`, `freeze_for_lm` with a frozen table): kernels only.
  * `freeze_for_alignment` (projector trainable, ViT + LLM frozen — the reference's ALIGNMENT stage): the tower runs
    on the kernels without a graph, the projector runs as `autograd_ops.ProjectorFunction` (forward + backward on the
    tcgen05 GEMM), the splice is differentiable (backward = row gather).
  * anything that needs a gradient this path does not produce (ViT parameters, the LLM embedding table) RAISES —
    nothing is detached silently.
"""
from __future__ import annotations

import types
from typing import Any, Dict, Optional

import torch

from . import ops
from .configs import TowerConfig
from .h2d import HostStager
from .vit import ProjectorWeights, TowerRunner, TowerWeights

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
        """Reference `ImageModality` whose forward runs on the B200 kernels."""
        config_class = B200RefImageConfig
        preprocessor_class = ImageProcessor
        b200_forward_calls = 0        # instrumentation for the tests: how often each entry really ran
        b200_forward_into_calls = 0

        def _runner(self) -> TowerRunner:
            dev = self.feature_extractor.device
            if dev.type != "cuda":
                raise ops.MMError("meditron_clip_b200 needs its parameters on a CUDA device "
                                  "(sm_100a kernels only, no CPU fallback)")
            # re-pack when the parameters were (re)loaded or moved: data_ptr + version of one tower weight; the
            # projector (the part that trains) is keyed separately so an optimizer step does not re-fold the tower
            pre = self.feature_extractor.vision_model.pre_layrnorm.weight
            tkey = (str(dev), pre.data_ptr(), pre._version)
            pkey = tuple((p.data_ptr(), p._version) for p in self.projector.parameters())
            if getattr(self, "_b200_tkey", None) != tkey:
                tower = _tower_from_hf(self.feature_extractor.vision_model.config)
                self._b200_runner = TowerRunner(TowerWeights(tower, self.state_dict(), dev))
                self._b200_tkey, self._b200_pkey = tkey, pkey
            elif self._b200_pkey != pkey:
                self._b200_runner.w.proj = ProjectorWeights(self.state_dict(), dev)
                self._b200_pkey = pkey
            return self._b200_runner

        def _pixels(self, inputs, dev) -> torch.Tensor:
            if torch.is_tensor(inputs):
                return inputs.to(dev, non_blocking=True)
            inputs = list(inputs)
            if inputs and not inputs[0].is_cuda:      # the collator's list of per-image CPU tensors
                if getattr(self, "_b200_stager", None) is None or self._b200_stager.device != dev:
                    self._b200_stager = HostStager(dev)
                return self._b200_stager.stack_upload(inputs)
            return torch.stack(inputs, dim=0)

        def _release(self):
            """after the kernels that read the uploaded pixels are enqueued: the device buffer may be refilled"""
            st = getattr(self, "_b200_stager", None)
            if st is not None:
                st.uploader.release()

        def _grad_mode(self) -> str:
            """'none' | 'projector' — raises for gradients this path cannot produce."""
            if not torch.is_grad_enabled():
                return "none"
            if any(p.requires_grad for p in self.feature_extractor.parameters()):
                raise NotImplementedError(
                    "meditron_clip_b200: the image tower has trainable parameters and autograd is on, but the ViT "
                    "backward is not part of the B200 path (SURVEY.md §8(f)#2).  Use freeze_modality_embedder() / "
                    "freeze_for_alignment() (projector-only training), or torch.no_grad() for inference.")
            return "projector" if any(p.requires_grad for p in self.projector.parameters()) else "none"

        def forward(self, inputs) -> torch.Tensor:                       # image_modality.py:130-137
            mode = self._grad_mode()
            type(self).b200_forward_calls += 1
            with torch.no_grad():
                r = self._runner()
                px = self._pixels(inputs, r.w.device)
                if mode == "none":
                    y = r.forward(px)
                    self._release()
                    return y
                feats = r.features(px).clone()        # the runner's feature buffer is reused by the next call
                self._release()
            from .autograd_ops import projector_forward
            y = projector_forward(feats, self.projector)
            return y.view(px.shape[0], r.cfg.num_patches, -1)

        def forward_into(self, inputs, embedded_tokens: torch.Tensor, dest_rows: torch.Tensor) -> torch.Tensor:
            """forward + splice fused into the last projector GEMM (no-grad path only)."""
            if self._grad_mode() != "none":
                raise RuntimeError("forward_into is the inference path; call forward() when the projector trains")
            type(self).b200_forward_into_calls += 1
            with torch.no_grad():
                r = self._runner()
                px = self._pixels(inputs, r.w.device)
                y = r.project(r.features(px), out=embedded_tokens, out_row_map=dest_rows)
                self._release()
                return y

    cls = AutoModality.register(MODEL_TYPE)(B200RefImageModality)
    _registered[MODEL_TYPE] = cls
    return cls


@torch.no_grad()
def b200_generate(model, batch: Dict[str, Any], max_new_tokens: int = 512, temperature: float = 0.1,
                  do_sample: bool = True, sync_every: int = 16, output_attentions: bool = True, **kwargs) -> torch.Tensor:
    """`MultiModalModelForCausalLM.generate` (model.py:528-640) without its per-token host round trips (SURVEY §8(f)
    rank 4).  Same arguments, same return value (`[batch, n_generated]` int64 on the host), same stopping rule; the LLM
    call itself is the reference's (`self.model(inputs_embeds=..., past_key_values=..., use_cache=True)`).  Differences,
    all outside the arithmetic:
      * the sampled / arg-max token, the finished mask and the generated ids stay ON THE DEVICE (the reference calls
        `.cpu()` on every token, samples row by row in a Python loop, and patches finished rows on the host, :611-628);
      * "all rows finished" is tested on the host only every `sync_every` steps; the tokens generated past the step
        where the reference would have stopped are cut off afterwards, so the result is the same;
      * the next token's embedding row comes from `mm_embed_gather`, the attention mask is one preallocated buffer.
    Greedy decoding is token-for-token identical to the reference; with `do_sample` one batched `torch.multinomial`
    replaces the per-row calls (same distribution, different draw order from the generator).
    `output_attentions=True` is what the reference passes to the LLM (:598; it makes HF use its eager attention); keep
    it for bit-identical logits, pass False to let the LLM use SDPA."""
    input_ids = batch["input_ids"]
    pmi = batch["processed_multimodal_inputs"]
    temperature = max(temperature, 1e-6)
    device = model.model.device
    B = input_ids.shape[0]
    eos = int(model.config.eos_token_idx)
    emb = model.embed_modalities_with_text(input_ids if hasattr(model, "b200_state") else input_ids.to(device), pmi)
    table = model.model.get_input_embeddings().weight
    attn0 = batch["attention_mask"].to(device)
    position_ids = batch["position_ids"].to(device)
    S = attn0.shape[1]
    mask = torch.ones(B, S + max_new_tokens, dtype=attn0.dtype, device=device)
    mask[:, :S] = attn0
    tokens = torch.full((max_new_tokens, B), eos, dtype=torch.long, device=device)
    finished = torch.zeros(B, dtype=torch.bool, device=device)
    all_done_at = torch.full((1,), max_new_tokens, dtype=torch.long, device=device)   # first step after which all rows are done
    past = None
    n = 0
    for i in range(max_new_tokens):
        if i > 0:
            position_ids = torch.full((B, 1), S + i - 1, dtype=torch.long, device=device)          # :585-588
        out = model.model(inputs_embeds=emb, attention_mask=mask[:, :S + i], position_ids=position_ids,
                          past_key_values=past, use_cache=True, output_attentions=output_attentions)
        past = out.past_key_values
        probs = torch.softmax(out.logits[:, -1, :] / temperature, dim=-1)                           # :605-607
        nxt = torch.multinomial(probs, 1).view(B) if do_sample else torch.argmax(probs, dim=-1)
        nxt = torch.where(finished, torch.full_like(nxt, eos), nxt)                                 # :620-622
        tokens[i] = nxt
        finished = finished | (nxt == eos)                                                          # :627-628
        all_done_at = torch.minimum(all_done_at, torch.where(finished.all(), i, max_new_tokens).view(1))
        n = i + 1
        if (i + 1) % sync_every == 0 and bool(finished.all()):                                      # :630 (amortised)
            break
        emb = ops.embed_gather(table.detach(), nxt.view(B, 1))                                      # :634-635
    stop = min(n, int(all_done_at.item()) + 1)
    return tokens[:stop].transpose(1, 0).cpu()


def _resolve_attachment_idx(attachment_token_idx, tokenizer, attachment_token) -> Optional[int]:
    if attachment_token_idx is not None:
        return int(attachment_token_idx)
    if tokenizer is not None and attachment_token is not None:
        idx = tokenizer.convert_tokens_to_ids(attachment_token)      # prompt_tokenizers.py:43
        if idx is None or idx == getattr(tokenizer, "unk_token_id", None):
            raise ValueError(f"attachment token {attachment_token!r} is not in the tokenizer's vocabulary")
        return int(idx)
    return None


def patch_embed_modalities_with_text(model, attachment_token_idx: Optional[int] = None, tokenizer=None,
                                     attachment_token: Optional[str] = None, positions_from: str = "device",
                                     eager_checks: bool = True):
    """Replace `model.embed_modalities_with_text` (model.py:410-446) of a reference MultiModalModelForCausalLM
    instance: embedding lookup by `mm_embed_gather`, positions by the device prefix sum (or the collator's vectors),
    splice fused into the projector's last GEMM epilogue when the modality offers `forward_into`.

    The reference model config does not store the attachment token id (it lives in the tokenizer,
    prompt_tokenizers.py:43), so `positions_from="device"` needs it here: pass `attachment_token_idx`, or `tokenizer`
    + `attachment_token`.  Without it this function RAISES rather than silently using the collator path.

    Errors (B3): with `eager_checks` and host-resident `input_ids` (what the collator produces) the attachment count and
    the id range are verified on the host BEFORE any launch and raise like the reference (RuntimeError / IndexError).
    For device-resident ids the check is lazy: call `model.b200_check()` (one sync) before trusting the result; the
    status word is cleared by every check, also when it raises.
    """
    if positions_from not in ("device", "collator"):
        raise ValueError(f"positions_from must be 'device' or 'collator', got {positions_from!r}")
    att = _resolve_attachment_idx(attachment_token_idx, tokenizer, attachment_token)
    if positions_from == "device" and att is None:
        raise ValueError("positions_from='device' needs the attachment token id: pass attachment_token_idx=, or "
                         "tokenizer= and attachment_token= (the reference config does not carry it)")
    state: Dict[str, Any] = {"status": None, "fused_calls": 0, "device_position_calls": 0, "calls": 0}

    def embed_modalities_with_text(self, input_ids, processed_multimodal_inputs):
        emb = self.model.get_input_embeddings()
        table = emb.weight
        dev = table.device
        if torch.is_grad_enabled() and table.requires_grad:
            raise NotImplementedError(
                "meditron_clip_b200: the LLM embedding table requires grad and autograd is on; the B200 splice path "
                "does not produce that gradient (it would be dropped).  Freeze the LLM (freeze_for_alignment), run "
                "under torch.no_grad(), or use the reference embed_modalities_with_text for this training mode.")
        stacks = processed_multimodal_inputs["stacked"]
        B, S = input_ids.shape
        H = table.shape[1]
        use_dev = positions_from == "device" and len(stacks) == 1
        if eager_checks and not input_ids.is_cuda:
            if input_ids.numel() and (int(input_ids.min()) < 0 or int(input_ids.max()) >= table.shape[0]):
                raise IndexError("index out of range in self")                       # nn.Embedding, model.py:433
            if use_dev:
                name, stack = next(iter(stacks.items()))
                want = len(stack) * self._get_modality_by_name(name)._num_patches_per_entry
                have = int((input_ids == att).sum())
                if want != have:                                                     # index_put_, model.py:441
                    raise RuntimeError(f"shape mismatch: value tensor of shape [{want}, {H}] cannot be broadcast to "
                                       f"indexing result of shape [{have}, {H}]")
        if state.get("stager") is None or state["stager"].device != dev:
            from .h2d import HostTensorStager
            state["stager"] = HostTensorStager(dev)
        ids = state["stager"].upload(input_ids)             # pageable host ids -> pinned staging -> async H2D
        if state["status"] is None or state["status"].device != dev:
            state["status"] = ops.new_status(dev)
        st = state["status"]
        state["calls"] += 1
        with torch.no_grad():
            embedded = ops.embed_gather(table.detach(), ids, status=st)             # table is frozen here (checked)
        for modality_name, stack in stacks.items():
            modality = self._get_modality_by_name(modality_name)
            trains = torch.is_grad_enabled() and any(p.requires_grad for p in modality.parameters())
            if use_dev:
                state["device_position_calls"] += 1
                n_rows = len(stack) * modality._num_patches_per_entry
                with torch.no_grad():
                    dest_rows, _, _ = ops.attach_positions(ids, att, expected_rows=n_rows, want_src_of_pos=False,
                                                           status=st)
                if not trains and hasattr(modality, "forward_into") and embedded.dtype == torch.bfloat16:
                    state["fused_calls"] += 1
                    modality.forward_into(stack, embedded, dest_rows)
                    continue
                proj = modality(stack)
                src = proj.reshape(-1, H).to(embedded.dtype)
                if proj.requires_grad:
                    from .autograd_ops import splice_rows
                    embedded = splice_rows(embedded.view(B * S, H), src, dest_rows, st).view(B, S, H)
                else:
                    ops.splice_scatter(embedded.view(B * S, H), src.contiguous(), dest_rows, status=st)
            else:
                proj = modality(stack)
                bi = state["stager"].upload(processed_multimodal_inputs["batch_idx"][modality_name])
                tr = state["stager"].upload(processed_multimodal_inputs["token_range"][modality_name])
                src = proj.reshape(-1, H).to(embedded.dtype)
                if proj.requires_grad:
                    from .autograd_ops import splice_rows
                    rows = (bi * S + tr).to(torch.int32)
                    embedded = splice_rows(embedded.view(B * S, H), src, rows, st).view(B, S, H)
                else:
                    ops.splice_scatter_idx(embedded, src.contiguous(), bi, tr, status=st)
        return embedded

    def b200_check():
        """Raise what the reference would have raised for the batches run since the last check; always clears."""
        st = state["status"]
        if st is None:
            return
        try:
            ops.raise_for_status(st)
        finally:
            st.zero_()

    model.embed_modalities_with_text = types.MethodType(embed_modalities_with_text, model)
    model.b200_check = b200_check
    model.b200_state = state
    model.b200_generate = types.MethodType(b200_generate, model)    # device-side decode loop (same signature as generate)
    if not getattr(model, "_b200_generate_wrapped", False):
        ref_generate = model.generate

        def generate(*a, **kw):          # model.py:571 calls embed_modalities_with_text outside its own no_grad block
            with torch.no_grad():
                return ref_generate(*a, **kw)

        model.generate = generate
        model._b200_generate_wrapped = True
    return model
