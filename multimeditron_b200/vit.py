"""Host-side driver of the image tower + projector on the sm_100a kernels.

`TowerWeights` packs a reference state dict (feature_extractor.vision_model.* / projector.projection.{0,2,4}.*)
once into the layouts the kernels want (fused [3D,D] QKV weight, zero-padded [D,Kpad] patch weight, bf16).
`TowerRunner.forward` issues, per encoder layer, 5 kernel launches on the current stream:
    GEMM(qkv, LN1 folded) -> attention -> GEMM(out,+bias,+residual, emits row stats) -> GEMM(fc1, LN2 folded, act) ->
    GEMM(fc2,+bias,+residual, emits row stats)
(the two LayerNorms are folded into the consuming GEMMs: statistics come out of the producing GEMM's epilogue, the
normalisation is applied in the consumer's epilogue — mm_gemm_ln_bias_act; only layer 0's LN1 is a separate launch),
then the 3 projector GEMMs (bias+erf-GELU epilogues); the last one can scatter its rows straight into
inputs_embeds (splice fused into the epilogue).  No host sync anywhere on the path.

Reference: image_modality.py:130-137; HF:models/clip/modeling_clip.py:667-692; projectors/mlp.py:33-39.
"""
from __future__ import annotations

import os

import torch

from . import ops
from .configs import TowerConfig

VM = "feature_extractor.vision_model."


def _bf(t: torch.Tensor, device) -> torch.Tensor:
    return t.detach().to(device=device, dtype=torch.bfloat16).contiguous()


# MM_B200_FOLD_LN=0 keeps the separate LayerNorm launches (A/B switch for tools and tests)
FOLD_LN = os.environ.get("MM_B200_FOLD_LN", "1") != "0"


class TowerWeights:
    """Device-resident, kernel-ready copy of one modality's parameters (replicated per GPU)."""

    def __init__(self, cfg: TowerConfig, sd: dict, device, vm_prefix: str = VM, with_projector: bool = True,
                 proj_prefix: str = "projector."):
        self.cfg = cfg
        self.device = torch.device(device)
        D = cfg.hidden_size
        dev = self.device
        w = sd[vm_prefix + "embeddings.patch_embedding.weight"].reshape(D, -1)
        kpad = (cfg.patch_k + 63) // 64 * 64
        wp = torch.zeros(D, kpad, dtype=torch.bfloat16, device=dev)
        wp[:, : cfg.patch_k] = _bf(w, dev)
        self.patch_w = wp
        self.patch_b = _bf(sd[vm_prefix + "embeddings.patch_embedding.bias"], dev) if cfg.conv_bias else None
        self.pos = _bf(sd[vm_prefix + "embeddings.position_embedding.weight"], dev)
        self.cls = _bf(sd[vm_prefix + "embeddings.class_embedding"], dev) if cfg.has_cls else None
        if cfg.kind == "clip":
            self.pre_ln = (_bf(sd[vm_prefix + "pre_layrnorm.weight"], dev), _bf(sd[vm_prefix + "pre_layrnorm.bias"], dev))
        else:
            self.pre_ln = None
            self.post_ln = (_bf(sd[vm_prefix + "post_layernorm.weight"], dev),
                            _bf(sd[vm_prefix + "post_layernorm.bias"], dev))
        self.layers = []
        nh, hd = cfg.num_attention_heads, cfg.head_dim
        hs = ops.attn_head_stride(hd)   # head slot in the qkv buffer (72 -> 80: 8 zero rows per head)
        self.head_stride = hs

        def slot(t: torch.Tensor) -> torch.Tensor:
            """[nh*hd, ...] -> [nh*hs, ...] with zero rows padding every head to its slot."""
            if hs == hd:
                return t
            out = torch.zeros((nh, hs) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            out[:, :hd] = t.view((nh, hd) + tuple(t.shape[1:]))
            return out.view((nh * hs,) + tuple(t.shape[1:]))

        for i in range(cfg.num_hidden_layers):
            L = f"{vm_prefix}encoder.layers.{i}."
            g = lambda n: _bf(sd[L + n], dev)  # noqa: E731
            self.layers.append(dict(
                ln1=(g("layer_norm1.weight"), g("layer_norm1.bias")),
                wqkv=torch.cat([slot(g("self_attn.q_proj.weight")), slot(g("self_attn.k_proj.weight")),
                                slot(g("self_attn.v_proj.weight"))], 0).contiguous(),
                bqkv=torch.cat([slot(g("self_attn.q_proj.bias")), slot(g("self_attn.k_proj.bias")),
                                slot(g("self_attn.v_proj.bias"))], 0).contiguous(),
                wo=g("self_attn.out_proj.weight"), bo=g("self_attn.out_proj.bias"),
                ln2=(g("layer_norm2.weight"), g("layer_norm2.bias")),
                w1=g("mlp.fc1.weight"), b1=g("mlp.fc1.bias"),
                w2=g("mlp.fc2.weight"), b2=g("mlp.fc2.bias"),
            ))
            if FOLD_LN:
                # LayerNorm folded into the consuming GEMM (mm_gemm_ln_bias_act): zero-row-sum W'' = gamma*W, b' = b + W beta.
                # LN1 of layer 0 keeps the LayerNorm kernel (its input comes from pre-LN / patch-embed, no statistics).
                Lw = self.layers[-1]
                Lw["w1_f"], Lw["b1_f"] = ops.fold_layernorm(Lw["w1"], Lw["b1"], *Lw["ln2"])
                if i > 0:
                    Lw["wqkv_f"], Lw["bqkv_f"] = ops.fold_layernorm(Lw["wqkv"], Lw["bqkv"], *Lw["ln1"])
                    del Lw["wqkv"], Lw["bqkv"]
                del Lw["w1"], Lw["b1"]
        self.proj = None
        if with_projector:
            self.proj = ProjectorWeights(sd, dev, proj_prefix)


class ProjectorWeights:
    """MLPProjector parameters (projectors/mlp.py:33-39) in bf16 on device."""

    def __init__(self, sd: dict, device, prefix: str = "projector."):
        p = prefix + "projection."
        self.w0, self.b0 = _bf(sd[p + "0.weight"], device), _bf(sd[p + "0.bias"], device)
        self.w2, self.b2 = _bf(sd[p + "2.weight"], device), _bf(sd[p + "2.bias"], device)
        self.w4, self.b4 = _bf(sd[p + "4.weight"], device), _bf(sd[p + "4.bias"], device)
        self.hidden = self.w4.shape[0]


class TowerRunner:
    """Runs the tower (and projector) for batches of images; activation workspaces are cached per batch size."""

    def __init__(self, weights: TowerWeights, shared_ws: dict | None = None):
        """shared_ws: a dict shared by several runners of the SAME tower shape (MoE experts run one after the other
        on one stream, so they can reuse one set of activation buffers)."""
        self.w = weights
        self.cfg = weights.cfg
        self._ws = shared_ws if shared_ws is not None else {}

    # -- workspaces ----------------------------------------------------------------------------------------
    def _workspace(self, B: int):
        # keyed by (batch, CUDA stream): two streams driving the same runner get separate activation buffers
        key = (B, torch.cuda.current_stream(self.w.device).cuda_stream)
        ws = self._ws.get(key)
        if ws is None:
            c, dev = self.cfg, self.w.device
            M, D, I = B * c.num_tokens, c.hidden_size, c.intermediate_size
            e = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=dev)  # noqa: E731
            ws = dict(stats=torch.zeros(M, ops.row_stats_slots(D), 2, dtype=torch.float32, device=dev), x=e(M, D), h=e(M, D), qkv=e(M, 3 * c.num_attention_heads * self.w.head_stride), attn=e(M, D), mlp=e(M, I), feats=e(B * c.num_patches, D))
            if c.has_cls:  # GEMM row m (token) -> compact feature row, CLS rows dropped (image_modality.py:133)
                m = torch.arange(M, device=dev, dtype=torch.int64)
                t = m % c.num_tokens
                ws["drop_cls"] = torch.where(t == 0, torch.full_like(m, -1),
                                             (m // c.num_tokens) * c.num_patches + t - 1).to(torch.int32)
            for k in [k for k in self._ws if k[1] == key[1]]:
                del self._ws[k]   # keep only the latest batch size resident per stream
            self._ws[key] = ws
        return ws

    # -- tower -----------------------------------------------------------------------------------------------
    def features(self, pixels: torch.Tensor, out: torch.Tensor | None = None,
                 out_row_map: torch.Tensor | None = None) -> torch.Tensor:
        """pixels [B,3,img,img] (fp32 or bf16, CUDA) -> encoder features [B*P, D] bf16 (CLS dropped).
        With (out, out_row_map): feature row (b*P+p) is written to out.view(-1,D)[out_row_map[b*P+p]] instead
        (int32 map over the B*P compact rows; used by the MoE stacks)."""
        c, w = self.cfg, self.w
        if pixels.shape[-1] != c.image_size or pixels.shape[-2] != c.image_size:
            # same exception class / message as HF:models/clip/modeling_clip.py:204-207
            raise ValueError(f"Input image size ({pixels.shape[-2]}*{pixels.shape[-1]}) doesn't match model "
                             f"({c.image_size}*{c.image_size}).")
        B = pixels.shape[0]
        ws = self._workspace(B)
        x, h, qkv, attn, mlp, feats = ws["x"], ws["h"], ws["qkv"], ws["attn"], ws["mlp"], ws["feats"]
        if out is not None:
            assert out.dtype == torch.bfloat16 and out.is_contiguous() and out.shape[-1] == c.hidden_size
            if out_row_map is None:
                assert out.numel() == feats.numel()
                feats = out.view(feats.shape)
        T, D, nh, hd = c.num_tokens, c.hidden_size, c.num_attention_heads, c.head_dim
        act = ops.ACT_BY_NAME[c.hidden_act]
        ops.patch_embed(pixels.contiguous(), w.patch_w, w.patch_b, w.pos, w.cls, c.patch_size, out=x.view(B, T, D))
        if w.pre_ln is not None:
            ops.layernorm(x, w.pre_ln[0], w.pre_ln[1], c.layer_norm_eps, out=x)
        n_layers = len(w.layers)
        stats, eps = ws["stats"], c.layer_norm_eps
        for i, L in enumerate(w.layers):
            last = i == n_layers - 1
            if "wqkv_f" in L:   # LN1 folded: statistics of x came out of the previous layer's fc2 epilogue
                ops.gemm_ln_bias_act(x, L["wqkv_f"], L["bqkv_f"], out=qkv, ln_stats=stats, eps=eps)
            else:
                ops.layernorm(x, L["ln1"][0], L["ln1"][1], eps, out=h)
                ops.gemm_bias_act(h, L["wqkv"], L["bqkv"], out=qkv)
            ops.attention_fwd(qkv, B, T, nh, hd, hd ** -0.5, out=attn)
            if "w1_f" in L:     # out-proj emits the row statistics LN2 needs; fc1 applies LN2 in its epilogue
                ops.gemm_ln_bias_act(attn, L["wo"], L["bo"], residual=x, out=x, stats_out=stats)
                ops.gemm_ln_bias_act(x, L["w1_f"], L["b1_f"], act=act, out=mlp, ln_stats=stats, eps=eps)
                if not last:
                    ops.gemm_ln_bias_act(mlp, L["w2"], L["b2"], residual=x, out=x, stats_out=stats)
                    continue
            else:
                ops.gemm_bias_act(attn, L["wo"], L["bo"], residual=x, out=x)
                ops.layernorm(x, L["ln2"][0], L["ln2"][1], eps, out=h)
                ops.gemm_bias_act(h, L["w1"], L["b1"], act=act, out=mlp)
            if last and c.has_cls:
                # last fc2 writes the compact [B*P, D] feature matrix directly (CLS rows never stored)
                rmap, dst = ws["drop_cls"], feats
                if out is not None and out_row_map is not None:   # compose: token row -> compact row -> caller's row
                    rmap = self._compose(ws, out_row_map)
                    dst = out.view(-1, D)
                ops.gemm_bias_act(mlp, L["w2"], L["b2"], residual=x, out=dst, out_row_map=rmap)
            else:
                ops.gemm_bias_act(mlp, L["w2"], L["b2"], residual=x, out=x)
        if not c.has_cls:
            if out is not None and out_row_map is not None:
                raise NotImplementedError("row-mapped feature output needs a CLS tower (MoE experts are CLIP)")
            ops.layernorm(x, w.post_ln[0], w.post_ln[1], c.layer_norm_eps, out=feats)
        return out if (out is not None and out_row_map is not None) else feats

    @staticmethod
    def _compose(ws, out_row_map: torch.Tensor) -> torch.Tensor:
        """token-row map = out_row_map[drop_cls[m]] (with -1 kept); cached per map TENSOR: the cache entry keeps the map
        alive (so its address cannot be handed to another tensor while the entry exists) and records its version
        counter (an in-place edit of the map invalidates the entry)."""
        key = ("compose", out_row_map.data_ptr(), out_row_map.numel())
        hit = ws.get(key)
        if hit is not None and hit[0] is out_row_map and hit[1] == out_row_map._version:
            return hit[2]
        dc = ws["drop_cls"].long()
        m = torch.where(dc >= 0, out_row_map.long()[dc.clamp(min=0)], dc).to(torch.int32)
        ws[key] = (out_row_map, out_row_map._version, m)
        return m

    # -- projector ---------------------------------------------------------------------------------------------
    def project(self, feats: torch.Tensor, proj: "ProjectorWeights | None" = None, out: torch.Tensor | None = None,
                out_row_map: torch.Tensor | None = None) -> torch.Tensor:
        """MLPProjector on [M', D] features.  With (out, out_row_map) the last GEMM's epilogue scatters each row
        to out.view(-1,H)[out_row_map[m]] (splice fused; model.py:441-444)."""
        p = proj or self.w.proj
        a = ops.gemm_bias_act(feats, p.w0, p.b0, act=ops.ACT_GELU_ERF)
        b = ops.gemm_bias_act(a, p.w2, p.b2, act=ops.ACT_GELU_ERF)
        if out is None:
            return ops.gemm_bias_act(b, p.w4, p.b4)
        assert out.dtype == torch.bfloat16 and out.is_contiguous()
        ops.gemm_bias_act(b, p.w4, p.b4, out=out.view(-1, p.hidden), out_row_map=out_row_map)
        return out

    def forward(self, pixels: torch.Tensor) -> torch.Tensor:
        """image_modality.py:130-137 -> [B, P, H] bf16."""
        B = pixels.shape[0]
        y = self.project(self.features(pixels))
        return y.view(B, self.cfg.num_patches, -1)


class GroupedTowerRunner:
    """E same-shaped towers (the MoE modalities' experts, image_modality_moe.py:160-162) on STACKED activations: every
    encoder GEMM of a layer is ONE `mm_gemm_grouped` launch for all experts (rows stacked expert-major, weights stacked
    along their rows), attention ONE launch over E*B images.  5 experts x 150 tiles of a D x D GEMM fill 10.1 waves of
    CTA pairs instead of 5 x 2.03 (the third wave of every per-expert launch was 3 % full).  Needs B*T % 256 == 0 (no row
    tile may straddle two experts); `eligible()` says so and the caller keeps the per-expert path otherwise.  Results are
    bit-identical to the per-expert runners (same kernels, same operands, same k order)."""

    def __init__(self, weights: list):
        assert len(weights) > 1 and all(w.cfg == weights[0].cfg for w in weights) and FOLD_LN
        self.ws_list, self.cfg, self.E = weights, weights[0].cfg, len(weights)
        self.device = weights[0].device

        def cat(key, i):
            """Stack one parameter of all experts along its rows ONCE and make every expert's own entry a view (row
            slice) of the stack: the per-expert runners (the fallback when `eligible()` says no) and the grouped path
            then read the same storage, instead of the experts' weights being resident twice."""
            parts = [w.layers[i][key] for w in weights]
            stacked = torch.cat(parts, 0).contiguous()
            n = parts[0].shape[0]
            for e, w in enumerate(weights):
                w.layers[i][key] = stacked[e * n:(e + 1) * n]      # contiguous, 16-byte aligned (n*K*2 B per expert)
            return stacked

        self.layers = []
        for i in range(self.cfg.num_hidden_layers):
            L = {k: cat(k, i) for k in ("wo", "bo", "w2", "b2", "w1_f", "b1_f")}
            for k in (("wqkv_f", "bqkv_f") if i > 0 else ("wqkv", "bqkv")):
                L[k] = cat(k, i)
            self.layers.append(L)
        self._ws = {}

    def eligible(self, B: int) -> bool:
        return (B * self.cfg.num_tokens) % 256 == 0 and self.cfg.has_cls

    def _workspace(self, B: int):
        key = (B, torch.cuda.current_stream(self.device).cuda_stream)
        ws = self._ws.get(key)
        if ws is None:
            c, dev, E = self.cfg, self.device, self.E
            M, D, I = E * B * c.num_tokens, c.hidden_size, c.intermediate_size
            e = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=dev)  # noqa: E731
            ws = dict(stats=torch.zeros(M, ops.row_stats_slots(D), 2, dtype=torch.float32, device=dev), x=e(M, D), h=e(M, D),
                      qkv=e(M, 3 * c.num_attention_heads * self.ws_list[0].head_stride), attn=e(M, D), mlp=e(M, I))
            m = torch.arange(M, device=dev, dtype=torch.int64)
            t = m % c.num_tokens
            ws["drop_cls"] = torch.where(t == 0, torch.full_like(m, -1),
                                         (m // c.num_tokens) * c.num_patches + t - 1).to(torch.int32)
            for k in [k for k in self._ws if k[1] == key[1]]:
                del self._ws[k]
            self._ws[key] = ws
        return ws

    def features(self, pixels: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """pixels [B,3,img,img] -> out [E, B*P, D] bf16 (every expert's CLS-dropped features, expert-major)."""
        c, E = self.cfg, self.E
        B = pixels.shape[0]
        if pixels.shape[-1] != c.image_size or pixels.shape[-2] != c.image_size:
            raise ValueError(f"Input image size ({pixels.shape[-2]}*{pixels.shape[-1]}) doesn't match model "
                             f"({c.image_size}*{c.image_size}).")
        assert self.eligible(B) and out.is_contiguous() and out.numel() == E * B * c.num_patches * c.hidden_size
        ws = self._workspace(B)
        x, h, qkv, attn, mlp, stats = ws["x"], ws["h"], ws["qkv"], ws["attn"], ws["mlp"], ws["stats"]
        T, D, nh, hd, eps = c.num_tokens, c.hidden_size, c.num_attention_heads, c.head_dim, c.layer_norm_eps
        Mg = B * T
        act = ops.ACT_BY_NAME[c.hidden_act]
        px = pixels.contiguous()
        for e, w in enumerate(self.ws_list):       # patch embed + pre-LN + layer-0 LN1: per-expert parameters on row slices
            xe = x[e * Mg:(e + 1) * Mg]
            ops.patch_embed(px, w.patch_w, w.patch_b, w.pos, w.cls, c.patch_size, out=xe.view(B, T, D))
            ops.layernorm(xe, w.pre_ln[0], w.pre_ln[1], eps, out=xe)
            ops.layernorm(xe, w.layers[0]["ln1"][0], w.layers[0]["ln1"][1], eps, out=h[e * Mg:(e + 1) * Mg])
        n_layers = len(self.layers)
        for i, L in enumerate(self.layers):
            if i == 0:
                ops.gemm_grouped(h, L["wqkv"], L["bqkv"], E, out=qkv)
            else:
                ops.gemm_grouped(x, L["wqkv_f"], L["bqkv_f"], E, out=qkv, ln_stats=stats, eps=eps)
            ops.attention_fwd(qkv, E * B, T, nh, hd, hd ** -0.5, out=attn)
            ops.gemm_grouped(attn, L["wo"], L["bo"], E, residual=x, out=x, stats_out=stats)
            ops.gemm_grouped(x, L["w1_f"], L["b1_f"], E, act=act, out=mlp, ln_stats=stats, eps=eps)
            if i < n_layers - 1:
                ops.gemm_grouped(mlp, L["w2"], L["b2"], E, residual=x, out=x, stats_out=stats)
            else:   # last fc2 writes the compact [E*B*P, D] feature matrix directly (CLS rows never stored)
                ops.gemm_grouped(mlp, L["w2"], L["b2"], E, residual=x, out=out.view(-1, D), out_row_map=ws["drop_cls"])
        return out
