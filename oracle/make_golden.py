"""Generate tests/golden/*.npz from the UNMODIFIED reference run in this container (TEST INFRASTRUCTURE).

    python -m oracle.make_golden            # needs /root/reference; writes tests/golden/

What is pinned:
  modality_<tower>.npz   reference ImageModality.forward (HF CLIPVisionTransformer + reference MLPProjector) on
                         seeded pixels/weights: fp32 run and pure-bf16 run (regime (i), §8 C5), plus taps for `tiny`
  splice.npz             reference MultiModalModelForCausalLM.embed_modalities_with_text (model.py:410-446) on
                         seeded cases: right/left padding, 0/1/4 images per row, image at position 0 / at the end,
                         fp32<->bf16 casts
  token_range.npz        reference PromptTokenizer.compute_token_range (prompt_tokenizers.py:255-286) + the collator's
                         index expansion statements (data_loader.py:136-141)
Weights/pixels are regenerated from (config, seed) by multimeditron_b200.synth at test time, so only outputs are stored.
The script also asserts that oracle/*.py reproduces the reference bit-for-bit on this CPU.
"""
from __future__ import annotations

import os
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from multimeditron_b200 import synth  # noqa: E402
from multimeditron_b200.configs import TOWERS  # noqa: E402
from oracle import ref_shim, splice_oracle, vit_oracle  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
MODALITY_CASES = {  # tower -> (llm hidden, batch, weight seed, pixel seed)
    "tiny": (64, 4, 11, 3),
    "clip_b32": (512, 8, 0, 0),   # BASELINE.json configs[0]
}


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy()


def build_reference_modality(ref, cfg, hidden, sd):
    from transformers import CLIPConfig, CLIPImageProcessor, CLIPModel
    tmp = tempfile.mkdtemp(prefix="clip_ckpt_")
    vc = dict(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
              num_hidden_layers=cfg.num_hidden_layers, num_attention_heads=cfg.num_attention_heads,
              image_size=cfg.image_size, patch_size=cfg.patch_size, layer_norm_eps=cfg.layer_norm_eps,
              hidden_act=cfg.hidden_act)
    tc = dict(hidden_size=32, intermediate_size=64, num_hidden_layers=1, num_attention_heads=2, vocab_size=64,
              max_position_embeddings=16)
    CLIPModel(CLIPConfig(vision_config=vc, text_config=tc, projection_dim=32)).save_pretrained(tmp)
    CLIPImageProcessor(size={"shortest_edge": cfg.image_size},
                       crop_size={"height": cfg.image_size, "width": cfg.image_size}).save_pretrained(tmp)
    from multimeditron.model.modalities.image_modality import ImageConfig, ImageModality
    m = ImageModality(ImageConfig(hidden_size=hidden, clip_name=tmp))
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(not k.startswith(("feature_extractor.vision_model", "projector")) for k in missing), missing
    return m.eval()


@torch.no_grad()
def gen_modality(ref):
    for name, (hidden, batch, wseed, pseed) in MODALITY_CASES.items():
        cfg = TOWERS[name]
        sd32 = synth.make_modality_state_dict(cfg, hidden, seed=wseed)
        pixels = synth.make_pixels(cfg, batch, seed=pseed)
        inputs = [pixels[i] for i in range(batch)]
        out = {}
        m = build_reference_modality(ref, cfg, hidden, sd32).float()
        # ---- fp32 reference run + module-boundary taps via forward hooks
        taps_ref = {}
        vm = m.feature_extractor.vision_model
        hooks = [vm.embeddings.register_forward_hook(lambda _m, _i, o: taps_ref.__setitem__("embeddings", o)),
                 vm.encoder.layers[0].register_forward_hook(lambda _m, _i, o: taps_ref.__setitem__("layer0", o)),
                 vm.encoder.layers[-1].register_forward_hook(
                     lambda _m, _i, o: taps_ref.__setitem__(f"layer{cfg.num_hidden_layers - 1}", o))]
        y32 = m(inputs)
        for h in hooks:
            h.remove()
        taps = {}
        o32 = vit_oracle.image_modality_forward(sd32, cfg, inputs, taps)
        assert torch.equal(y32, o32), f"{name}: oracle != reference (fp32)"
        for k, v in taps_ref.items():
            v = v[0] if isinstance(v, tuple) else v
            assert torch.equal(v, taps[k]), f"{name}: tap {k} differs"
        out["out_fp32"] = y32.numpy()
        if name == "tiny":
            for k in ("embeddings", "layer0", f"layer{cfg.num_hidden_layers - 1}", "features"):
                out["tap_" + k] = taps[k].numpy()
        # ---- pure-bf16 module run (regime (i): training dtype, every op rounds to bf16)
        mb = m.to(torch.bfloat16)
        yb = mb(inputs)
        sdb = {k: v.to(torch.bfloat16) for k, v in sd32.items()}
        ob = vit_oracle.image_modality_forward(sdb, cfg, inputs)
        assert torch.equal(yb, ob), f"{name}: oracle != reference (bf16)"
        out["out_bf16_bits"] = bf16_bits(yb)
        err = (yb.float() - y32).abs().max().item()
        print(f"[modality {name}] out {tuple(y32.shape)} absmax {y32.abs().max():.4f} bf16-vs-fp32 maxabs {err:.3e}")
        np.savez_compressed(os.path.join(GOLD, f"modality_{name}.npz"), hidden=hidden, batch=batch, wseed=wseed,
                            pseed=pseed, **out)


@torch.no_grad()
def gen_siglip():
    """Row A9 (no reference class): HF SiglipVisionModel.last_hidden_state (all tokens, post-LN applied,
    HF:models/siglip/modeling_siglip.py:604-625) + the reference MLPProjector (projectors/mlp.py:33-59)."""
    from transformers import SiglipVisionConfig, SiglipVisionModel
    from multimeditron.model.projectors.mlp import MLPProjector
    name, hidden, batch, wseed, pseed = "tiny_siglip", 96, 3, 5, 9
    cfg = TOWERS[name]
    sd32 = synth.make_modality_state_dict(cfg, hidden, seed=wseed)
    pixels = synth.make_pixels(cfg, batch, seed=pseed)
    hf = SiglipVisionModel(SiglipVisionConfig(
        hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size, num_hidden_layers=cfg.num_hidden_layers,
        num_attention_heads=cfg.num_attention_heads, image_size=cfg.image_size, patch_size=cfg.patch_size,
        layer_norm_eps=cfg.layer_norm_eps, hidden_act=cfg.hidden_act)).eval()
    pre = "feature_extractor."
    missing, unexpected = hf.load_state_dict({k[len(pre):]: v for k, v in sd32.items() if k.startswith(pre)}, strict=False)
    assert not unexpected and all(k.startswith("vision_model.head.") for k in missing), (missing, unexpected)
    proj = MLPProjector(cfg.hidden_size, hidden, dtype=torch.float32)
    proj.load_state_dict({k[len("projector."):]: v for k, v in sd32.items() if k.startswith("projector.")})
    out = {}
    for dt, tag in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        hf, proj = hf.to(dt), proj.to(dt)
        y = proj(hf(pixels.to(dt)).last_hidden_state)
        sd = {k: v.to(dt) for k, v in sd32.items()}
        o = vit_oracle.image_modality_forward(sd, cfg, [pixels[i] for i in range(batch)])
        assert torch.equal(y, o), f"siglip oracle != HF ({tag})"
        out[tag] = y
    err = (out["bf16"].float() - out["fp32"]).abs().max().item()
    print(f"[modality {name}] out {tuple(out['fp32'].shape)} absmax {out['fp32'].abs().max():.4f} bf16-vs-fp32 maxabs {err:.3e}")
    np.savez_compressed(os.path.join(GOLD, f"modality_{name}.npz"), hidden=hidden, batch=batch, wseed=wseed, pseed=pseed,
                        out_fp32=out["fp32"].numpy(), out_bf16_bits=bf16_bits(out["bf16"]))


PREPROCESS_CASES = [  # (H, W, size, square) raw uint8 images, seeded
    (37, 53, 56, False), (300, 400, 224, False), (640, 480, 224, False), (224, 224, 224, False), (1000, 333, 224, False),
    (90, 120, 224, False), (225, 224, 224, False), (500, 700, 384, True), (100, 80, 384, True), (336, 336, 336, False),
]


def gen_preprocess():
    """`ImageProcessor.process` (image_modality.py:96) = HF CLIPImageProcessor / SiglipImageProcessor (transformers
    5.5.0, torchvision backend) on seeded uint8 PIL images; oracle/preprocess_oracle.py must match it bit for bit."""
    from PIL import Image
    from transformers import CLIPImageProcessor, SiglipImageProcessor
    from oracle import preprocess_oracle as po
    out = {"cases": np.array(PREPROCESS_CASES, np.int64)}
    for i, (h, w, size, square) in enumerate(PREPROCESS_CASES):
        img = synth.make_raw_image(i, h, w)   # noise, every third one a smooth ramp
        if square:
            p = SiglipImageProcessor(size={"height": size, "width": size})
            mean = std = (0.5, 0.5, 0.5)
        else:
            p = CLIPImageProcessor(size={"shortest_edge": size}, crop_size={"height": size, "width": size})
            mean, std = po.OPENAI_CLIP_MEAN, po.OPENAI_CLIP_STD
        ref = p(images=Image.fromarray(img), return_tensors="pt")["pixel_values"][0].numpy()
        got = po.clip_preprocess(img, size, mean, std, square=bool(square))
        assert np.array_equal(ref.view(np.int32), got.view(np.int32)), f"preprocess oracle != HF for case {i}"
        # store the uint8 stage (compact) — the fp32 value is an exact function of it
        m, s = po.fused_mean_std(mean, std)
        u8 = np.rint(ref * s[:, None, None] + m[:, None, None]).astype(np.uint8)
        assert np.array_equal(((u8.astype(np.float32) - m[:, None, None]) / s[:, None, None]).view(np.int32), ref.view(np.int32))
        import hashlib
        out[f"sha_{i}"] = np.frombuffer(hashlib.sha256(ref.tobytes()).digest(), np.uint8)   # digest of the fp32 tensor
        if size <= 56:
            out[f"u8_{i}"] = u8
    np.savez_compressed(os.path.join(GOLD, "preprocess.npz"), **out)
    print(f"[preprocess] {len(PREPROCESS_CASES)} cases pinned against the HF processors")


SPLICE_CASES = [
    # name, batch, seq, images/row, tokens/image, hidden, padding, min_len, table dtype, proj dtype
    ("right_1img", 3, 40, 1, 9, 32, "right", 24, "float32", "float32"),
    ("left_4img", 4, 96, 4, 9, 32, "left", 60, "float32", "float32"),
    ("noimg_rows", 5, 48, 2, 4, 16, "right", 20, "float32", "float32"),
    ("img_at_edges", 2, 22, 2, 9, 16, "right", 22, "float32", "float32"),
    ("bf16_table_f32_proj", 3, 40, 2, 9, 32, "left", 30, "bfloat16", "float32"),
    ("f32_table_bf16_proj", 3, 40, 2, 9, 32, "right", 30, "float32", "bfloat16"),
    ("bf16_both", 4, 64, 3, 9, 64, "right", 40, "bfloat16", "bfloat16"),
]
VOCAB, ATT, START, END, PAD = 97, 3, 4, 5, 0


def splice_case_inputs(case):
    name, B, S, n_img, tpi, H, side, min_len, tdt, pdt = case
    seed = sum(map(ord, name))
    ids, mask = synth.make_sequences(B, S, n_img, tpi, VOCAB, ATT, START, END, PAD, seed=seed, padding_side=side,
                                     min_len=min_len)
    if name == "noimg_rows":      # rows 1 and 3 carry no image at all
        for b in (1, 3):
            row = ids[b]
            row[(row == ATT) | (row == START) | (row == END)] = 11
    if name == "img_at_edges":    # image block at position 0 of row 0 and flush at the end of row 1
        blk = torch.tensor([START] + [ATT] * tpi + [END])
        ids[0] = torch.cat([blk, torch.full((S - len(blk),), 13)])
        ids[1] = torch.cat([torch.full((S - len(blk),), 17), blk])
    gen = torch.Generator().manual_seed(seed)
    table = (torch.randn(VOCAB, H, generator=gen) * 0.02).to(getattr(torch, tdt))
    n_mod = int((ids == ATT).sum())
    proj = torch.randn(n_mod // tpi, tpi, H, generator=gen).to(getattr(torch, pdt))
    return ids, table, proj, tpi


@torch.no_grad()
def gen_splice(ref):
    Model = ref.model.MultiModalModelForCausalLM
    pt_cls = ref.prompt_tokenizers
    tok_cls = [getattr(pt_cls, n) for n in dir(pt_cls) if hasattr(getattr(pt_cls, n), "compute_token_range")][0]
    out, tr_out = {}, {}
    for case in SPLICE_CASES:
        name = case[0]
        ids, table, proj, tpi = splice_case_inputs(case)
        # --- indices exactly as the reference builds them: compute_token_range per row + collator expansion
        fake_tok = types.SimpleNamespace(attachment_token_idx=ATT, get_num_embeddings=lambda m: m["num_embeddings"])
        ranges = []
        for b in range(ids.shape[0]):
            n_img = int((ids[b] == ATT).sum()) // tpi
            mods = [{"num_embeddings": tpi, "type": "image"} for _ in range(n_img)]
            ranges.append(tok_cls.compute_token_range(fake_tok, ids[b], mods))
        multi = [(b, tr) for b, rs in enumerate(ranges) for tr in rs]
        if multi:   # data_loader.py:136-139, same statements
            batch_idx, token_range = zip(*multi)
            bidx = torch.tensor(batch_idx).repeat_interleave(torch.tensor([tr[1] - tr[0] for tr in token_range]))
            trng = torch.cat([torch.tensor(range(tr[0], tr[1])) for tr in token_range])
        else:
            bidx = trng = torch.zeros(0, dtype=torch.int64)
        # oracle restatement must agree with the reference's statements
        o_ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [tpi] * (int((ids[b] == ATT).sum()) // tpi), ATT)
                    for b in range(ids.shape[0])]
        assert o_ranges == [[tuple(map(int, r)) for r in rs] for rs in ranges], name
        ob, ot = splice_oracle.collator_indices(o_ranges)
        assert np.array_equal(ob, bidx.numpy()) and np.array_equal(ot, trng.numpy()), name
        dest, _ = splice_oracle.attach_positions(ids.numpy(), ATT)
        assert np.array_equal(dest, (bidx * ids.shape[1] + trng).numpy().astype(np.int32)), name
        # --- the reference's own embed_modalities_with_text on a minimal stand-in for `self`
        emb = torch.nn.Embedding(VOCAB, table.shape[1], _weight=table.clone())
        fake_self = types.SimpleNamespace(
            model=types.SimpleNamespace(get_input_embeddings=lambda: emb),
            _get_modality_by_name=lambda n: (lambda stack: proj))
        pmi = {"stacked": {"image": [None]}, "batch_idx": {"image": bidx}, "token_range": {"image": trng}}
        y = Model.embed_modalities_with_text(fake_self, ids, pmi)
        o = splice_oracle.embed_modalities_with_text(table, ids, bidx, trng, proj)
        assert torch.equal(y, o), name
        out[name + "__ids"] = ids.numpy()
        out[name + "__bidx"] = bidx.numpy()
        out[name + "__trng"] = trng.numpy()
        out[name + "__out"] = bf16_bits(y) if y.dtype == torch.bfloat16 else y.numpy()
        tr_out[name + "__ranges"] = np.array([(b, s, e) for b, rs in enumerate(ranges) for (s, e) in rs],
                                             dtype=np.int64).reshape(-1, 3)
        print(f"[splice {name}] ids {tuple(ids.shape)} n_mod {bidx.numel()} out {tuple(y.shape)} {y.dtype}")
    np.savez_compressed(os.path.join(GOLD, "splice.npz"), **out)
    np.savez_compressed(os.path.join(GOLD, "token_range.npz"), **tr_out)


POSITION_CASES = [  # name, batch, seq, images/row, grid, padding, min_len
    ("right_1img", 3, 40, 1, 3, "right", 24), ("left_3img", 4, 64, 3, 3, "left", 40), ("noimg", 2, 20, 0, 3, "right", 9),
    ("edges", 2, 22, 2, 3, "right", 22), ("grid4_left", 3, 60, 2, 4, "left", 45),
]


def gen_position_ids(ref):
    """Runs the UNMODIFIED second half of DataCollatorForMultimodal.torch_call (data_loader.py:120-195); only the
    tokenisation front half (SamplePreprocessor / loaders, not on the path) is replaced by pre-built features."""
    from unittest import mock
    dl = ref.data_loader
    out = {}
    for name, B, S, n_img, g, side, min_len in POSITION_CASES:
        tpi = g * g
        ids, mask = synth.make_sequences(B, S, n_img, tpi, VOCAB, ATT, START, END, PAD, seed=len(name), padding_side=side,
                                         min_len=min_len)
        if name == "edges":   # image block at position 0 of row 0 and flush at the end of row 1 (the `s > 0` quirk, :178)
            blk = torch.tensor([START] + [ATT] * tpi + [END])
            ids[0] = torch.cat([blk[1:], torch.full((S - len(blk) + 1,), 13)])        # attachment run starts at t = 0
            ids[1] = torch.cat([torch.full((S - 2 * len(blk),), 17), blk, blk])
            mask[:] = True
        ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [tpi] * (int((ids[b] == ATT).sum()) // tpi), ATT)
                  for b in range(B)]
        res = {}
        for use_2d in (False, True):
            feats = []
            for b in range(B):
                mods = []
                for (s0, e0) in ranges[b]:
                    pm = {"type": "image", "value": torch.zeros(1), "token_range": (s0, e0), "num_embeddings": tpi}
                    if use_2d:   # what ImageProcessor.process emits with use_2d_position_ids (image_modality.py:99-108)
                        pm["position_ids"] = torch.stack(torch.meshgrid(
                            torch.arange(g, dtype=torch.long), torch.arange(g, dtype=torch.long), indexing="ij"),
                            dim=-1).reshape(tpi, 2)
                    mods.append(pm)
                feats.append({"input_ids": ids[b].clone(), "labels": ids[b].clone(),
                              "attention_mask": mask[b].long(), "modalities": mods})
            fake_sp = mock.MagicMock()
            fake_sp.return_value.process_modality_to_tensor.side_effect = lambda raw: raw
            fake_sp.return_value.tokenize.side_effect = lambda processed, add_generation_prompt=False: processed
            with mock.patch.object(dl, "SamplePreprocessor", fake_sp), \
                    mock.patch.object(dl.BaseModalityLoader, "load_modalities", staticmethod(lambda f, loaders: f)):
                coll = dl.DataCollatorForMultimodal(tokenizer=None, modality_processors={}, modality_loaders={},
                                                    attachment_token="<att>", chat_template=None,
                                                    use_2d_position_ids=use_2d)
                batch = coll.torch_call(feats)
            got = batch["position_ids"].numpy()
            want = splice_oracle.position_ids(mask.numpy(), ranges, g, use_2d)
            assert got.dtype == np.int64 and np.array_equal(got, want), f"position_ids oracle != reference ({name}, 2d={use_2d})"
            res["2d" if use_2d else "1d"] = got
            if not use_2d:   # the collator's index build, again from the real statements
                ob, ot = splice_oracle.collator_indices(ranges)
                pmi = batch["processed_multimodal_inputs"]
                if ob.size:
                    assert np.array_equal(pmi["batch_idx"]["image"].numpy(), ob)
                    assert np.array_equal(pmi["token_range"]["image"].numpy(), ot)
        out[name + "__ids"], out[name + "__mask"] = ids.numpy(), mask.numpy()
        out[name + "__grid"] = np.int64(g)
        out[name + "__pos1d"], out[name + "__pos2d"] = res["1d"], res["2d"]
        print(f"[position_ids {name}] {tuple(ids.shape)} grid {g}: 1-D and 2-D pinned against the reference collator")
    np.savez_compressed(os.path.join(GOLD, "position_ids.npz"), **out)


MOE_CASES = [  # name, per-expert projector, fusion, experts, hidden, batch, weight seed, pixel seed
    ("moe_shared_avg", False, "weighted_average", 3, 64, 3, 21, 5),
    ("moe_shared_append", False, "sequence_append", 3, 64, 2, 22, 6),
    ("moe_pep_avg", True, "weighted_average", 3, 64, 3, 23, 7),
    ("moe_pep_append", True, "sequence_append", 3, 64, 2, 24, 8),
    # CLIP ViT-B/32-sized experts (the cfg5 expert shape: D 768, 12 layers, 12 heads, T = 50) -> hidden 512
    ("moe_shared_avg_b32", False, "weighted_average", 3, 512, 2, 31, 9, "clip_b32"),
    ("moe_pep_avg_b32", True, "weighted_average", 3, 512, 2, 33, 10, "clip_b32"),
    # cross_attn fusion (attention.py:48-101): tiny (dim 128 / 64, 2 heads -> head_dim 64 / 32), generalist_idx 1 and the
    # default -1 (where the raw-index comparison makes every expert a "specialist"); ViT-B/32 experts: 8 heads x 96
    ("moe_shared_xattn", False, "cross_attn", 3, 64, 3, 41, 11, "tiny", 1, 2),
    ("moe_pep_xattn", True, "cross_attn", 3, 64, 2, 42, 12, "tiny", -1, 2),
    ("moe_shared_xattn_b32", False, "cross_attn", 3, 512, 2, 43, 13, "clip_b32", -1, 8),
]
MOE_CLASS_ORDER = [2, 0, 1]   # gating class c is expert MOE_CLASS_ORDER[c]  (non-identity on purpose)


def save_tiny_clip(cfg):
    from transformers import CLIPConfig, CLIPImageProcessor, CLIPModel
    tmp = tempfile.mkdtemp(prefix="clip_ckpt_")
    vc = dict(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
              num_hidden_layers=cfg.num_hidden_layers, num_attention_heads=cfg.num_attention_heads,
              image_size=cfg.image_size, patch_size=cfg.patch_size, layer_norm_eps=cfg.layer_norm_eps,
              hidden_act=cfg.hidden_act)
    tc = dict(hidden_size=32, intermediate_size=64, num_hidden_layers=1, num_attention_heads=2, vocab_size=64,
              max_position_embeddings=16)
    CLIPModel(CLIPConfig(vision_config=vc, text_config=tc, projection_dim=32)).save_pretrained(tmp)
    CLIPImageProcessor(size={"shortest_edge": cfg.image_size},
                       crop_size={"height": cfg.image_size, "width": cfg.image_size}).save_pretrained(tmp)
    return tmp


@torch.no_grad()
def gen_moe(ref):
    """Reference MOEImageModality / MOEImageModalityPEP (all experts + gating-weighted fusion) on the `tiny` tower."""
    from multimeditron.model.modalities.image_modality_moe import MOEImageConfig, MOEImageModality
    from multimeditron.model.modalities.image_modality_moe_pep import MOEImageConfigPEP, MOEImageModalityPEP
    from multimeditron.model.modalities.moe.gating import GatingNetwork, GatingNetworkConfig
    only = os.environ.get("MOE_ONLY")
    for case in MOE_CASES:
        name, pep, fusion, E, hidden, batch, wseed, pseed = case[:8]
        tower = case[8] if len(case) > 8 else "tiny"
        gen_idx, xheads = (case[9], case[10]) if len(case) > 10 else (-1, 8)
        xattn = fusion == "cross_attn"
        if only and only not in name:
            continue
        cfg = TOWERS[tower]
        dirs = [save_tiny_clip(cfg) for _ in range(E)]
        gdir = tempfile.mkdtemp(prefix="gating_")
        torch.manual_seed(wseed)
        GatingNetwork(GatingNetworkConfig(num_classes=E, top_k=1, image_processor_path=dirs[0],
                                          class_names=[dirs[i] for i in MOE_CLASS_ORDER])).save_pretrained(gdir)
        Cfg, Mod = (MOEImageConfigPEP, MOEImageModalityPEP) if pep else (MOEImageConfig, MOEImageModality)
        m = Mod(Cfg(hidden_size=hidden, expert_clip_names=dirs, image_processor=dirs[0], gating_path=gdir,
                    top_k_experts=E if fusion == "sequence_append" else 1, fusion_method=fusion,
                    generalist_idx=gen_idx, cross_attn_heads=xheads)).eval()
        m = m.float()   # the MoE projectors are hard-defaulted to bf16 (projectors/mlp.py:16): cast BEFORE loading
        sd32 = synth.make_moe_state_dict([cfg] * E, hidden, pep, seed=wseed, cross_attn=xattn)
        missing, unexpected = m.load_state_dict(sd32, strict=False)
        assert not unexpected, unexpected
        assert all(k.startswith("gating_network.") for k in missing), missing
        # a random-init ResNet-50 gates ~uniformly, which would hide the weight->expert alignment: spread the logits
        m.gating_network.resnet.fc.bias.data = torch.tensor([1.5, -0.5, 0.25][:E]) + 0.1 * torch.arange(E)
        pixels = synth.make_pixels(cfg, batch, seed=pseed)
        inputs = [pixels[i] for i in range(batch)]
        _, _, w32 = m.gating_network(pixels)
        y32 = m(inputs)
        perm = m._gating_to_expert_perm.tolist()
        assert perm == MOE_CLASS_ORDER
        o32 = vit_oracle.moe_modality_forward(sd32, [cfg] * E, inputs, w32, perm, fusion, pep, gen_idx, xheads)
        assert torch.equal(y32, o32), f"{name}: oracle != reference (fp32)"
        mb = m.to(torch.bfloat16)
        _, _, wb = mb.gating_network(pixels.to(torch.bfloat16))
        yb = mb([p.to(torch.bfloat16) for p in inputs])
        sdb = {k: v.to(torch.bfloat16) for k, v in sd32.items()}
        ob = vit_oracle.moe_modality_forward(sdb, [cfg] * E, [p.to(torch.bfloat16) for p in inputs], wb, perm, fusion, pep,
                                             gen_idx, xheads)
        assert torch.equal(yb, ob), f"{name}: oracle != reference (bf16)"
        # the reference's own bf16 error with the SAME gating weights as the fp32 run (isolates tower/projector rounding)
        ob_w32 = vit_oracle.moe_modality_forward(sdb, [cfg] * E, [p.to(torch.bfloat16) for p in inputs],
                                                 w32.to(torch.bfloat16), perm, fusion, pep, gen_idx, xheads)
        err = (ob_w32.float() - y32).abs().max().item()
        print(f"[{name}] out {tuple(y32.shape)} absmax {y32.abs().max():.4f} bf16-vs-fp32 maxabs {err:.3e} "
              f"weights {w32[0].tolist()}")
        np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), pep=pep, fusion=fusion, experts=E, hidden=hidden,
                            batch=batch, wseed=wseed, pseed=pseed, perm=np.array(perm), weights32=w32.numpy(), tower=tower,
                            generalist_idx=gen_idx, cross_attn_heads=xheads,
                            out_fp32=y32.numpy(), out_bf16_w32_bits=bf16_bits(ob_w32))


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    ref = ref_shim.import_reference()
    only = sys.argv[1:]
    if not only or "splice" in only:
        gen_splice(ref)
    if not only or "modality" in only:
        gen_modality(ref)
    if not only or "siglip" in only:
        gen_siglip()
    if not only or "position_ids" in only:
        gen_position_ids(ref)
    if not only or "preprocess" in only:
        gen_preprocess()
    if not only or "moe" in only:
        gen_moe(ref)
    print("golden written to", GOLD)


if __name__ == "__main__":
    main()
