"""-m "not gpu": the oracle against the golden vectors produced by the reference itself (oracle/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import synth
from multimeditron_b200.configs import TOWERS
from oracle import splice_oracle, vit_oracle
from oracle.make_golden import ATT, SPLICE_CASES, splice_case_inputs


def _bits(t):
    return t.contiguous().view(torch.int16).numpy()


@pytest.mark.parametrize("name", ["tiny", "clip_b32", "tiny_siglip"])
def test_modality_oracle_matches_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"modality_{name}.npz"))
    cfg = TOWERS[name]
    hidden, batch = int(g["hidden"]), int(g["batch"])
    sd = synth.make_modality_state_dict(cfg, hidden, seed=int(g["wseed"]))
    px = synth.make_pixels(cfg, batch, seed=int(g["pseed"]))
    taps = {}
    with torch.no_grad():
        y = vit_oracle.image_modality_forward(sd, cfg, [px[i] for i in range(batch)], taps)
    ref = torch.from_numpy(g["out_fp32"])
    # same torch build + same ops => bit-identical; allow last-bit slack for a different host CPU's GEMM blocking
    assert y.shape == ref.shape
    assert torch.allclose(y, ref, rtol=0, atol=2e-5 * ref.abs().max().item())
    for k in g.files:
        if k.startswith("tap_"):
            t = torch.from_numpy(g[k])
            assert torch.allclose(taps[k[4:]], t, rtol=0, atol=2e-5 * t.abs().max().item()), k
    if name == "tiny":  # bf16 regime as well (cheap)
        sdb = {k: v.to(torch.bfloat16) for k, v in sd.items()}
        with torch.no_grad():
            yb = vit_oracle.image_modality_forward(sdb, cfg, [px[i] for i in range(batch)])
        refb = torch.from_numpy(g["out_bf16_bits"]).view(torch.bfloat16)
        assert (yb.float() - refb.float()).abs().max().item() <= 2e-2 * refb.float().abs().max().item()


@pytest.mark.parametrize("case", SPLICE_CASES, ids=[c[0] for c in SPLICE_CASES])
def test_splice_oracle_bit_exact_vs_reference_golden(golden_dir, case):
    g = np.load(os.path.join(golden_dir, "splice.npz"))
    tr = np.load(os.path.join(golden_dir, "token_range.npz"))
    name = case[0]
    ids, table, proj, tpi = splice_case_inputs(case)
    assert np.array_equal(ids.numpy(), g[name + "__ids"])
    # index construction: compute_token_range + collator expansion
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [tpi] * (int((ids[b] == ATT).sum()) // tpi), ATT)
              for b in range(ids.shape[0])]
    flat = np.array([(b, s, e) for b, rs in enumerate(ranges) for (s, e) in rs], dtype=np.int64).reshape(-1, 3)
    assert np.array_equal(flat, tr[name + "__ranges"])
    bidx, trng = splice_oracle.collator_indices(ranges)
    assert np.array_equal(bidx, g[name + "__bidx"]) and np.array_equal(trng, g[name + "__trng"])
    # prefix-sum positions == collator indices (SURVEY.md §3.4 invariant)
    dest, sop = splice_oracle.attach_positions(ids.numpy(), ATT)
    assert np.array_equal(dest.astype(np.int64), bidx * ids.shape[1] + trng)
    assert np.array_equal(np.flatnonzero(sop >= 0), dest)
    out = splice_oracle.embed_modalities_with_text(table, ids, torch.from_numpy(bidx), torch.from_numpy(trng), proj)
    want = g[name + "__out"]
    got = _bits(out) if out.dtype == torch.bfloat16 else out.numpy()
    assert got.dtype == want.dtype and np.array_equal(got, want)


def test_splice_shape_mismatch_raises_like_reference():
    ids, table, proj, tpi = splice_case_inputs(SPLICE_CASES[0])
    dest, _ = splice_oracle.attach_positions(ids.numpy(), ATT)
    bidx = torch.from_numpy(dest.astype(np.int64) // ids.shape[1])
    trng = torch.from_numpy(dest.astype(np.int64) % ids.shape[1])
    with pytest.raises(RuntimeError):
        splice_oracle.embed_modalities_with_text(table, ids, bidx[:-1], trng[:-1], proj)


@pytest.mark.needs_reference
def test_oracle_against_live_hf_tower():
    """Cross-check of the restatement against the installed transformers (the third-party home of the arithmetic)."""
    from transformers import CLIPVisionConfig, CLIPVisionModel
    cfg = TOWERS["tiny"]
    hf = CLIPVisionModel(CLIPVisionConfig(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
                                          num_hidden_layers=cfg.num_hidden_layers,
                                          num_attention_heads=cfg.num_attention_heads, image_size=cfg.image_size,
                                          patch_size=cfg.patch_size)).eval()
    sd = synth.make_modality_state_dict(cfg, 64, seed=3)
    vis = {k[len("feature_extractor."):]: v for k, v in sd.items() if k.startswith("feature_extractor.")}
    hf.load_state_dict(vis, strict=True)
    px = synth.make_pixels(cfg, 2, seed=9)
    with torch.no_grad():
        ref = hf(px).last_hidden_state
        got = vit_oracle.vision_last_hidden_state(sd, cfg, px)
    assert torch.equal(ref, got)


def test_c_oracle_agrees_with_numpy_oracle():
    from oracle import c_oracle
    rng = np.random.default_rng(0)
    ids = rng.integers(5, 50, size=(7, 33)).astype(np.int64)
    ids[rng.random(ids.shape) < 0.3] = 3
    d_np, s_np = splice_oracle.attach_positions(ids, 3)
    d_c, s_c = c_oracle.attach_positions(ids, 3)
    assert np.array_equal(d_np, d_c) and np.array_equal(s_np, s_c)
    table = rng.standard_normal((50, 16)).astype(np.float32)
    src = rng.standard_normal((d_np.shape[0], 16)).astype(np.float32)
    out = c_oracle.embed_splice(table, ids, src, d_np)
    want = splice_oracle.embed_modalities_with_text(
        torch.from_numpy(table), torch.from_numpy(ids), torch.from_numpy(d_np.astype(np.int64) // 33),
        torch.from_numpy(d_np.astype(np.int64) % 33), torch.from_numpy(src)).numpy().reshape(-1, 16)
    assert np.array_equal(out, want)


MOE_NAMES = ["moe_shared_avg", "moe_shared_append", "moe_pep_avg", "moe_pep_append", "moe_shared_avg_b32", "moe_pep_avg_b32",
             "moe_shared_xattn", "moe_pep_xattn", "moe_shared_xattn_b32"]


@pytest.mark.parametrize("name", MOE_NAMES)
def test_moe_oracle_matches_reference_golden(golden_dir, name):
    """MOEImageModality / MOEImageModalityPEP run by oracle/make_golden.py (all experts, gating-weighted fusion, the
    PEP 'no permutation' quirk with a non-identity class order)."""
    g = np.load(os.path.join(golden_dir, f"{name}.npz"))
    cfg = TOWERS[str(g["tower"]) if "tower" in g.files else "tiny"]
    E, hidden, batch, pep = int(g["experts"]), int(g["hidden"]), int(g["batch"]), bool(g["pep"])
    xattn = str(g["fusion"]) == "cross_attn"
    gi, xh = (int(g["generalist_idx"]), int(g["cross_attn_heads"])) if "generalist_idx" in g.files else (-1, 8)
    sd = synth.make_moe_state_dict([cfg] * E, hidden, pep, seed=int(g["wseed"]), cross_attn=xattn)
    px = synth.make_pixels(cfg, batch, seed=int(g["pseed"]))
    perm = g["perm"].tolist()
    assert perm != sorted(perm)
    with torch.no_grad():
        y = vit_oracle.moe_modality_forward(sd, [cfg] * E, px, torch.from_numpy(g["weights32"]), perm, str(g["fusion"]), pep,
                                            gi, xh)
        ref = torch.from_numpy(g["out_fp32"])
        assert y.shape == ref.shape
        assert torch.allclose(y, ref, rtol=0, atol=2e-5 * ref.abs().max().item())
        if str(g["fusion"]) == "weighted_average" and not pep:   # the alignment matters: identity must NOT match
            wrong = vit_oracle.moe_modality_forward(sd, [cfg] * E, px, torch.from_numpy(g["weights32"]),
                                                    list(range(E)), "weighted_average", pep)
            assert not torch.allclose(wrong, ref, rtol=0, atol=1e-3 * ref.abs().max().item())
