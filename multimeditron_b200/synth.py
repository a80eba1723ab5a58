"""Deterministic synthetic weights / inputs with the reference's state-dict key names.

There is no network for checkpoints, so tests and bench.py use random-init towers.  Per SURVEY.md
§8 C4 the reference's own random init leaves every bias 0 and LayerNorm at (1, 0), which would never
exercise the fused bias/LN paths, so biases ~ N(0,0.02), gamma ~ 1+N(0,0.1), beta ~ N(0,0.1).
Values are drawn on the CPU generator in a fixed key order, so a (config, seed) pair names the same
tensors on every machine with this torch build.
"""
from __future__ import annotations

import torch

from .configs import TowerConfig

VM = "feature_extractor.vision_model."


def tower_keys(cfg: TowerConfig):
    """(key, shape, kind) in canonical order; kind in {w, b, g(amma), beta}."""
    D, I, p = cfg.hidden_size, cfg.intermediate_size, cfg.patch_size
    out = []
    if cfg.has_cls:
        out.append((VM + "embeddings.class_embedding", (D,), "w"))
    out.append((VM + "embeddings.patch_embedding.weight", (D, 3, p, p), "w"))
    if cfg.conv_bias:
        out.append((VM + "embeddings.patch_embedding.bias", (D,), "b"))
    out.append((VM + "embeddings.position_embedding.weight", (cfg.num_tokens, D), "w"))
    if cfg.kind == "clip":
        out += [(VM + "pre_layrnorm.weight", (D,), "g"), (VM + "pre_layrnorm.bias", (D,), "beta")]
    for i in range(cfg.num_hidden_layers):
        L = f"{VM}encoder.layers.{i}."
        out += [(L + "layer_norm1.weight", (D,), "g"), (L + "layer_norm1.bias", (D,), "beta")]
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            out += [(L + f"self_attn.{n}.weight", (D, D), "w"), (L + f"self_attn.{n}.bias", (D,), "b")]
        out += [(L + "layer_norm2.weight", (D,), "g"), (L + "layer_norm2.bias", (D,), "beta")]
        out += [(L + "mlp.fc1.weight", (I, D), "w"), (L + "mlp.fc1.bias", (I,), "b")]
        out += [(L + "mlp.fc2.weight", (D, I), "w"), (L + "mlp.fc2.bias", (D,), "b")]
    out += [(VM + "post_layernorm.weight", (D,), "g"), (VM + "post_layernorm.bias", (D,), "beta")]
    return out


def projector_keys(D: int, H: int, prefix: str = "projector."):
    """MLPProjector: projection.{0,2,4} (src/multimeditron/model/projectors/mlp.py:33-39)."""
    return [
        (prefix + "projection.0.weight", (D, D), "w"), (prefix + "projection.0.bias", (D,), "b"),
        (prefix + "projection.2.weight", (H, D), "w"), (prefix + "projection.2.bias", (H,), "b"),
        (prefix + "projection.4.weight", (H, H), "w"), (prefix + "projection.4.bias", (H,), "b"),
    ]


def _draw(keys, gen, dtype, device):
    sd = {}
    for k, shape, kind in keys:
        t = torch.randn(shape, generator=gen, dtype=torch.float32)
        if kind in ("w", "b"):
            t = t * 0.02
        elif kind == "g":
            t = 1.0 + 0.1 * t
        else:
            t = 0.1 * t
        sd[k] = t.to(dtype).to(device)
    return sd


def make_modality_state_dict(cfg: TowerConfig, llm_hidden: int, seed: int = 0, dtype=torch.float32, device="cpu"):
    """State dict of one `meditron_clip`-style modality: feature_extractor.vision_model.* + projector.*"""
    gen = torch.Generator().manual_seed(seed)
    return _draw(tower_keys(cfg) + projector_keys(cfg.hidden_size, llm_hidden), gen, dtype, device)


def add_massive_activations(sd: dict, cfg: TowerConfig, seed: int = 0, frac: float = 0.03) -> dict:
    """Trained CLIP/SigLIP towers carry a few residual-stream channels whose magnitude is 50-200x the rest ("massive
    activations"); random-init towers do not, so the LayerNorm statistics (E[x^2]-mean^2 in the folded GEMMs) are never
    stressed.  This plants them: `frac` of the channels get a large constant offset (+-U(50,200)) and a x U(5,20) gain at
    the point where the residual stream is born (CLIP: pre_layrnorm gamma/beta; SigLIP: the position embedding)."""
    g = torch.Generator().manual_seed(seed)
    D = cfg.hidden_size
    n = max(1, int(round(frac * D)))
    ch = torch.randperm(D, generator=g)[:n]
    off = (50.0 + 150.0 * torch.rand(n, generator=g)) * (torch.randint(0, 2, (n,), generator=g) * 2 - 1)
    gain = 5.0 + 15.0 * torch.rand(n, generator=g)
    out = dict(sd)
    if cfg.kind == "clip":
        w, b = sd[VM + "pre_layrnorm.weight"].clone(), sd[VM + "pre_layrnorm.bias"].clone()
        w[ch] = (w[ch].float() * gain).to(w.dtype)
        b[ch] = (b[ch].float() + off).to(b.dtype)
        out[VM + "pre_layrnorm.weight"], out[VM + "pre_layrnorm.bias"] = w, b
    else:
        pos = sd[VM + "embeddings.position_embedding.weight"].clone()
        pos[:, ch] = (pos[:, ch].float() * gain + off).to(pos.dtype)
        out[VM + "embeddings.position_embedding.weight"] = pos
    return out


def make_moe_state_dict(towers, llm_hidden: int, per_expert_projector: bool, seed: int = 0, dtype=torch.float32,
                        device="cpu", cross_attn: bool = False):
    """State dict of a MoE modality (image_modality_moe.py / image_modality_moe_pep.py): `experts.<i>.*` holds the HF
    CLIPVisionTransformer of expert i, `projector.*` (shared) or `projectors.<i>.*` (PEP) the MLP projectors."""
    gen = torch.Generator().manual_seed(seed)
    keys = []
    for i, cfg in enumerate(towers):
        keys += [(f"experts.{i}." + k[len(VM):], shape, kind) for k, shape, kind in tower_keys(cfg)]
        if per_expert_projector:
            keys += projector_keys(cfg.hidden_size, llm_hidden, prefix=f"projectors.{i}.")
    if not per_expert_projector:
        keys += projector_keys(towers[0].hidden_size, llm_hidden, prefix="projector.")
    sd = _draw(keys, gen, dtype, device)
    if cross_attn:   # CrossAttention (attention.py:36-42): dim = experts' width (shared) or the LLM hidden size (PEP)
        C = llm_hidden if per_expert_projector else towers[0].hidden_size
        xk = []
        for n in ("q_proj", "k_proj", "v_proj", "proj"):
            xk += [(f"cross_attn.{n}.weight", (C, C), "w"), (f"cross_attn.{n}.bias", (C,), "x")]
        xs = _draw(xk, gen, torch.float32, "cpu")  # drawn AFTER the rest: the other tensors do not depend on the flag
        for n in ("q_proj", "k_proj"):             # N(0, 0.06): scores spread enough for a non-uniform softmax
            xs[f"cross_attn.{n}.weight"] *= 3.0
        sd.update({k: v.to(dtype).to(device) for k, v in xs.items()})
    return sd


def make_pixels(cfg: TowerConfig, batch: int, seed: int = 0, device="cpu"):
    """Post-normalisation CLIP pixels are ~N(0,1); fp32 like the processor output."""
    gen = torch.Generator().manual_seed(1000 + seed)
    return torch.randn(batch, 3, cfg.image_size, cfg.image_size, generator=gen, dtype=torch.float32).to(device)


def make_sequences(batch: int, seq_len: int, images_per_row: int, tokens_per_image: int, vocab: int, attach_id: int,
                   start_id: int, end_id: int, pad_id: int, seed: int = 0, padding_side: str = "right",
                   min_len: int | None = None):
    """input_ids [batch, seq_len] int64 in the reference's expanded form
    (<image_start> + N x attachment + <image_end> per image, prompt_tokenizers.py:288-306), images at seeded
    sorted offsets, remaining ids uniform in [0, vocab) excluding the special ids, ragged rows padded."""
    gen = torch.Generator().manual_seed(2000 + seed)
    special = {attach_id, start_id, end_id, pad_id}
    block = tokens_per_image + 2
    ids = torch.full((batch, seq_len), pad_id, dtype=torch.int64)
    mask = torch.zeros((batch, seq_len), dtype=torch.bool)
    for b in range(batch):
        lo = max(min_len if min_len is not None else seq_len, images_per_row * block)
        length = int(torch.randint(lo, seq_len + 1, (1,), generator=gen)) if lo < seq_len else seq_len
        n_text = length - images_per_row * block
        assert n_text >= 0, "sequence too short for the requested images"
        text = torch.randint(0, vocab, (n_text,), generator=gen)
        for s in special:  # remap collisions deterministically
            text[text == s] = (s + 7) % vocab if (s + 7) % vocab not in special else 0
        cuts = torch.sort(torch.randint(0, n_text + 1, (images_per_row,), generator=gen)).values.tolist()
        row, prev = [], 0
        img = torch.tensor([start_id] + [attach_id] * tokens_per_image + [end_id], dtype=torch.int64)
        for c in cuts:
            row += [text[prev:c], img]
            prev = c
        row.append(text[prev:])
        row = torch.cat(row)
        if padding_side == "right":
            ids[b, :length] = row
            mask[b, :length] = True
        else:
            ids[b, seq_len - length:] = row
            mask[b, seq_len - length:] = True
    return ids, mask


def make_raw_image(i: int, h: int, w: int):
    """Seeded raw uint8 [h,w,3] image (noise; every third one a smooth ramp) for the pre-processing tests/goldens."""
    import numpy as np
    img = np.random.default_rng(100 + i).integers(0, 256, (h, w, 3), dtype=np.uint8)
    if i % 3 == 0:
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([(yy * 255 // max(1, h - 1)), (xx * 255 // max(1, w - 1)), ((yy + xx) % 256)], -1).astype(np.uint8)
    return img


def load_mock_images(n: int):
    """The two real photographs of the reference's `mock_dataset/` (cat.jpg 1200x1200, EPFL_campus_2017.jpg 2048x1365),
    committed DOWNSCALED (Lanczos, lossless PNG) under tests/golden/mock/ because /root/reference does not travel to the
    GPU box; cycled to `n` images ("cycle the 2 images to make batch 8").  -> list of uint8 [H,W,3] tensors."""
    import glob
    import os

    import numpy as np
    from PIL import Image
    d = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "mock")
    files = sorted(glob.glob(os.path.join(d, "*.png")))
    if not files:
        raise FileNotFoundError(f"no mock images under {d}")
    imgs = [torch.from_numpy(np.array(Image.open(f).convert("RGB"), dtype=np.uint8)) for f in files]
    return [imgs[i % len(imgs)] for i in range(n)]
