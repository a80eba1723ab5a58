"""Torch restatement of the ViT + projector arithmetic of the hot path (TEST INFRASTRUCTURE).

The arithmetic lives in the third-party dependency `transformers` (unpinned in the reference's
pyproject.toml:25; version installed in this image and used for all pins: 5.5.0).  Every function cites
the HF / reference lines it restates and calls the same torch ops in the same order, so that on the same
device and dtype it is bit-identical to the reference modules (checked by oracle/make_golden.py and
tests/test_oracle_golden.py).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

VM = "feature_extractor.vision_model."


def _act(name: str, x: torch.Tensor) -> torch.Tensor:
    if name == "quick_gelu":          # HF:activations.py:117-123  input * sigmoid(1.702 * input)
        return x * torch.sigmoid(1.702 * x)
    if name == "gelu":                # nn.GELU() default (erf)
        return F.gelu(x)
    if name == "gelu_pytorch_tanh":   # HF:activations.py  GELUTanh -> F.gelu(approximate="tanh")
        return F.gelu(x, approximate="tanh")
    raise ValueError(name)


def vision_embeddings(sd, cfg, pixels: torch.Tensor, vm: str = VM) -> torch.Tensor:
    """HF:models/clip/modeling_clip.py:202-219 (CLIP) / models/siglip/modeling_siglip.py:157-187 (SigLIP)."""
    w = sd[vm + "embeddings.patch_embedding.weight"]
    if pixels.shape[-1] != cfg.image_size or pixels.shape[-2] != cfg.image_size:
        raise ValueError(
            f"Input image size ({pixels.shape[-2]}*{pixels.shape[-1]}) doesn't match model "
            f"({cfg.image_size}*{cfg.image_size}).")
    b = sd.get(vm + "embeddings.patch_embedding.bias") if cfg.conv_bias else None
    x = F.conv2d(pixels.to(dtype=w.dtype), w, b, stride=cfg.patch_size)      # :208-209
    x = x.flatten(2).transpose(1, 2)                                         # :210
    pos = sd[vm + "embeddings.position_embedding.weight"]
    if cfg.has_cls:
        cls = sd[vm + "embeddings.class_embedding"].expand(x.shape[0], 1, -1)  # :212
        x = torch.cat([cls, x], dim=1)                                          # :213
    return x + pos.unsqueeze(0)                                              # :217


def encoder_layer(sd, cfg, i: int, x: torch.Tensor, vm: str = VM) -> torch.Tensor:
    """HF:models/clip/modeling_clip.py:363-385 (layer), :300-336 (attention, SDPA), :347-351 (MLP)."""
    L = f"{vm}encoder.layers.{i}."
    D, nh, hd = cfg.hidden_size, cfg.num_attention_heads, cfg.head_dim
    B, T, _ = x.shape
    res = x
    h = F.layer_norm(x, (D,), sd[L + "layer_norm1.weight"], sd[L + "layer_norm1.bias"], cfg.layer_norm_eps)
    q = F.linear(h, sd[L + "self_attn.q_proj.weight"], sd[L + "self_attn.q_proj.bias"])
    k = F.linear(h, sd[L + "self_attn.k_proj.weight"], sd[L + "self_attn.k_proj.bias"])
    v = F.linear(h, sd[L + "self_attn.v_proj.weight"], sd[L + "self_attn.v_proj.bias"])
    q = q.view(B, T, nh, hd).transpose(1, 2)
    k = k.view(B, T, nh, hd).transpose(1, 2)
    v = v.view(B, T, nh, hd).transpose(1, 2)
    # HF:integrations/sdpa_attention.py:40 — no mask, dropout 0, scale = head_dim**-0.5, not causal
    a = F.scaled_dot_product_attention(q, k, v, attn_mask=None, dropout_p=0.0, scale=hd ** -0.5, is_causal=False)
    a = a.transpose(1, 2).reshape(B, T, D).contiguous()
    a = F.linear(a, sd[L + "self_attn.out_proj.weight"], sd[L + "self_attn.out_proj.bias"])
    x = res + a
    res = x
    h = F.layer_norm(x, (D,), sd[L + "layer_norm2.weight"], sd[L + "layer_norm2.bias"], cfg.layer_norm_eps)
    h = F.linear(h, sd[L + "mlp.fc1.weight"], sd[L + "mlp.fc1.bias"])
    h = _act(cfg.hidden_act, h)
    h = F.linear(h, sd[L + "mlp.fc2.weight"], sd[L + "mlp.fc2.bias"])
    return res + h


def vision_last_hidden_state(sd, cfg, pixels: torch.Tensor, taps: dict | None = None, vm: str = VM) -> torch.Tensor:
    """CLIP: HF:models/clip/modeling_clip.py:667-692 — pre_layrnorm, L layers, last_hidden_state WITHOUT
    post_layernorm.  SigLIP: HF:models/siglip/modeling_siglip.py:604-625 — no pre-LN, post_layernorm on all tokens."""
    D = cfg.hidden_size
    x = vision_embeddings(sd, cfg, pixels, vm)
    if taps is not None:
        taps["embeddings"] = x
    if cfg.kind == "clip":
        x = F.layer_norm(x, (D,), sd[vm + "pre_layrnorm.weight"], sd[vm + "pre_layrnorm.bias"], cfg.layer_norm_eps)
    for i in range(cfg.num_hidden_layers):
        x = encoder_layer(sd, cfg, i, x, vm)
        if taps is not None and (i == 0 or i == cfg.num_hidden_layers - 1):
            taps[f"layer{i}"] = x
    if cfg.kind == "siglip":
        x = F.layer_norm(x, (D,), sd[vm + "post_layernorm.weight"], sd[vm + "post_layernorm.bias"], cfg.layer_norm_eps)
    return x


def mlp_projector(sd, x: torch.Tensor, prefix: str = "projector.") -> torch.Tensor:
    """src/multimeditron/model/projectors/mlp.py:33-39,57: Linear, GELU(erf), Linear, GELU(erf), Linear."""
    p = prefix + "projection."
    x = F.linear(x, sd[p + "0.weight"], sd[p + "0.bias"])
    x = F.gelu(x)
    x = F.linear(x, sd[p + "2.weight"], sd[p + "2.bias"])
    x = F.gelu(x)
    return F.linear(x, sd[p + "4.weight"], sd[p + "4.bias"])


def image_modality_forward(sd, cfg, inputs, taps: dict | None = None) -> torch.Tensor:
    """src/multimeditron/model/modalities/image_modality.py:130-137:
    stack -> vision_model(...).last_hidden_state[:, 1:, :] -> projector.
    For SigLIP (no reference class) all tokens are kept."""
    pixels = torch.stack(list(inputs), dim=0) if not torch.is_tensor(inputs) else inputs
    feats = vision_last_hidden_state(sd, cfg, pixels, taps)
    if cfg.has_cls:
        feats = feats[:, 1:, :]
    if taps is not None:
        taps["features"] = feats
    return mlp_projector(sd, feats)


def moe_weighted_average(expert_feats: torch.Tensor, weights: torch.Tensor, perm=None) -> torch.Tensor:
    """image_modality_moe.py:171-177 (weights permuted to expert order) / image_modality_moe_pep.py:212-216
    (no permutation): stacked [B,E,P,D] * w[B,E,1,1] summed over E."""
    w = weights if perm is None else weights[:, perm]
    w = w.to(expert_feats.dtype).unsqueeze(-1).unsqueeze(-1)
    return (expert_feats * w).sum(dim=1)


def moe_modality_forward(sd, towers, inputs, weights: torch.Tensor, perm, fusion_method: str,
                         per_expert_projector: bool, generalist_idx: int = -1, cross_attn_heads: int = 8) -> torch.Tensor:
    """image_modality_moe.py:152-210 (shared projector: weights permuted :171-175, projector after fusion :207) and
    image_modality_moe_pep.py:191-249 (projector per expert :198-203, NO permutation :212-216).
    `weights` [B,E] = the gating network's softmax output (moe/gating.py:85-87) — an input here."""
    pixels = torch.stack(list(inputs), dim=0) if not torch.is_tensor(inputs) else inputs
    outs = []
    for i, cfg in enumerate(towers):
        f = vision_last_hidden_state(sd, cfg, pixels, vm=f"experts.{i}.")[:, 1:, :]
        if per_expert_projector:
            f = mlp_projector(sd, f, prefix=f"projectors.{i}.")
        outs.append(f)
    stacked = torch.stack(outs, dim=1)                                   # [B,E,P,*]
    if fusion_method == "sequence_append":
        fused = torch.flatten(stacked, start_dim=1, end_dim=2)
    elif fusion_method == "weighted_average":
        w = weights
        if not per_expert_projector:
            w = w.index_select(dim=-1, index=torch.as_tensor(perm, dtype=torch.long, device=w.device))
        fused = (stacked * w.unsqueeze(-1).unsqueeze(-1)).sum(dim=1)
    elif fusion_method == "cross_attn":
        fused = cross_attn_fusion(sd, stacked, weights, perm, generalist_idx, cross_attn_heads)
    else:
        raise ValueError(f"Unsupported fusion_method: {fusion_method}")
    return fused if per_expert_projector else mlp_projector(sd, fused, prefix="projector.")


def cross_attention(sd, x: torch.Tensor, experts, num_heads: int, prefix: str = "cross_attn.") -> torch.Tensor:
    """src/multimeditron/model/attention.py:48-101 in eval mode (both dropouts are the identity), no mask."""
    B, N_q, C = x.shape
    hd = C // num_heads
    context = torch.cat(experts, dim=1)                                               # :67
    N_kv = context.shape[1]
    q = F.linear(x, sd[prefix + "q_proj.weight"], sd[prefix + "q_proj.bias"])          # :71-73
    k = F.linear(context, sd[prefix + "k_proj.weight"], sd[prefix + "k_proj.bias"])
    v = F.linear(context, sd[prefix + "v_proj.weight"], sd[prefix + "v_proj.bias"])
    q = q.view(B, N_q, num_heads, hd).transpose(1, 2)                                  # :76-78
    k = k.view(B, N_kv, num_heads, hd).transpose(1, 2)
    v = v.view(B, N_kv, num_heads, hd).transpose(1, 2)
    attn = torch.matmul(q, k.transpose(-2, -1)) * (hd ** -0.5)                         # :81
    attn = torch.softmax(attn, dim=-1)                                                 # :91
    out = torch.matmul(attn, v).transpose(1, 2).reshape(B, N_q, C)                     # :95-98
    return F.linear(out, sd[prefix + "proj.weight"], sd[prefix + "proj.bias"])         # :99


def cross_attn_fusion(sd, stacked: torch.Tensor, weights: torch.Tensor, perm, generalist_idx: int, num_heads: int):
    """image_modality_moe.py:178-205 == image_modality_moe_pep.py:218-245: generalist queries over gate-weighted
    specialists.  `i != generalist_idx` compares the RAW index (with the default -1 every expert is a specialist)."""
    B, E, P, C = stacked.shape
    q = stacked[:, generalist_idx, :, :]
    spec = [i for i in range(E) if i != generalist_idx]
    w_all = weights.index_select(dim=-1, index=torch.as_tensor(perm, dtype=torch.long, device=weights.device))
    w_spec = torch.softmax(w_all[:, spec], dim=-1)
    ctx = [stacked[:, e, :, :] * w_spec[:, j].view(B, 1, 1) for j, e in enumerate(spec)]
    return cross_attention(sd, q, ctx, num_heads)
