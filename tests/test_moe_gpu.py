"""-m gpu: MoE image modalities through the plugin mirror vs the goldens produced by the reference
`MOEImageModality` / `MOEImageModalityPEP`.  Tolerance: max|Δ| vs the reference fp32 run <= 2 x the reference's own
bf16 error + 1e-4·range, and <= 2e-2 of range."""
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import synth
from multimeditron_b200.configs import TOWERS
from multimeditron_b200.modalities import AutoModality
from multimeditron_b200.model import B200MultiModalEmbedder
from oracle import splice_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"
MOE_NAMES = ["moe_shared_avg", "moe_shared_append", "moe_pep_avg", "moe_pep_append",
             "moe_shared_avg_b32", "moe_pep_avg_b32",       # *_b32: CLIP ViT-B/32-sized experts (the cfg5 expert shape)
             "moe_shared_xattn", "moe_pep_xattn", "moe_shared_xattn_b32"]   # cross_attn fusion (attention.py:48-101)


def _build(g):
    tower = str(g["tower"]) if "tower" in g.files else "tiny"
    cfg = TOWERS[tower]
    E, hidden, pep = int(g["experts"]), int(g["hidden"]), bool(g["pep"])
    names = [f"expert{i}" for i in range(E)]
    perm = g["perm"].tolist()
    mcfg = AutoModality.config_from_dict({
        "model_type": "moe_meditron_clip_pep_b200" if pep else "moe_meditron_clip_b200", "hidden_size": hidden,
        "expert_clip_names": names, "expert_towers": {n: tower for n in names}, "image_processor": tower,
        "fusion_method": str(g["fusion"]), "top_k_experts": E if str(g["fusion"]) == "sequence_append" else 1,
        "gating_class_names": [names[i] for i in perm],
        "generalist_idx": int(g["generalist_idx"]) if "generalist_idx" in g.files else -1,
        "cross_attn_heads": int(g["cross_attn_heads"]) if "cross_attn_heads" in g.files else 8})
    sd = synth.make_moe_state_dict([cfg] * E, hidden, pep, seed=int(g["wseed"]),
                                   cross_attn=str(g["fusion"]) == "cross_attn")
    w = torch.from_numpy(g["weights32"])
    m = AutoModality.from_config(mcfg).load_reference_state_dict(sd, DEV, gating=lambda px: w.to(px.device))
    return m, cfg


@pytest.mark.parametrize("name", MOE_NAMES)
def test_moe_forward_vs_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"{name}.npz"))
    m, cfg = _build(g)
    batch = int(g["batch"])
    px = synth.make_pixels(cfg, batch, seed=int(g["pseed"]))
    y = m([px[i] for i in range(batch)])
    ref32 = torch.from_numpy(g["out_fp32"])
    refbf = torch.from_numpy(g["out_bf16_w32_bits"]).view(torch.bfloat16).float()
    assert y.shape == ref32.shape and y.dtype == torch.bfloat16
    scale = ref32.abs().max().item()
    err = (y.float().cpu() - ref32).abs().max().item()
    ref_err = (refbf - ref32).abs().max().item()
    print(f"[{name}] max-abs {err:.3e} (reference bf16 run: {ref_err:.3e}), max-abs/absmax {err / scale:.3e}")
    assert err <= 2.0 * ref_err + 1e-4 * scale
    assert err / scale <= 2e-2


@pytest.mark.parametrize("name", ["moe_shared_append", "moe_pep_avg"])
def test_moe_splice_fused_equals_unfused(golden_dir, name):
    """forward_into (splice fused into the modality) must be bit-identical to forward + splice_scatter."""
    g = np.load(os.path.join(golden_dir, f"{name}.npz"))
    m, cfg = _build(g)
    B, H, V, S = int(g["batch"]), int(g["hidden"]), 300, 80
    rows = m.rows_per_entry
    att, start, end, pad = V - 4, V - 3, V - 2, V - 1
    ids, _ = synth.make_sequences(B, S, 1, rows, V - 4, att, start, end, pad, seed=3, padding_side="left", min_len=S - 6)
    px = synth.make_pixels(cfg, B, seed=int(g["pseed"]))
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [rows], att) for b in range(B)]
    bidx, trng = splice_oracle.collator_indices(ranges)
    pmi = {"stacked": {"image": [px[i] for i in range(B)]}, "batch_idx": {"image": torch.from_numpy(bidx)},
           "token_range": {"image": torch.from_numpy(trng)}}
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(1)) * 0.02).bfloat16().to(DEV)
    outs = []
    for fuse in (True, False):
        for pos in ("device", "collator"):
            e = B200MultiModalEmbedder(table, [m], att, fuse_splice=fuse)
            outs.append(e(ids, pmi, positions_from=pos))
            e.check()
    for o in outs[1:]:
        assert torch.equal(o, outs[0])
    proj = m(pmi["stacked"]["image"]).cpu()
    want = splice_oracle.embed_modalities_with_text(table.cpu(), ids, pmi["batch_idx"]["image"],
                                                    pmi["token_range"]["image"], proj)
    assert torch.equal(outs[0].cpu(), want)


def test_resnet50_gate_matches_the_reference_gating_network():
    """moe/gating.py:73-89: ResNet-50 logits -> top-k -> softmax.  The device gate (bf16, channels-last, BatchNorm folded
    into the convolutions) against the same torchvision network run unfused in fp32 (what GatingNetwork executes)."""
    pytest.importorskip("torchvision")
    from multimeditron_b200.gating import ResNet50Gate
    torch.manual_seed(3)
    ref = ResNet50Gate(5, top_k=2, device=DEV, dtype=torch.float32, fold_bn=False)
    for m in ref.resnet.modules():          # non-trivial running statistics / affine parameters
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.normal_(0, 0.1); m.running_var.uniform_(0.5, 1.5)
            m.weight.data.uniform_(0.5, 1.5); m.bias.data.normal_(0, 0.1)
    sd = {"gating_network.resnet." + k: v for k, v in ref.resnet.state_dict().items()}
    gate = ResNet50Gate(5, top_k=2, device=DEV, state_dict=sd, prefix="gating_network.resnet.")
    px = torch.randn(6, 3, 224, 224, device=DEV)
    l32, t32, w32 = ref(px)
    l, t, w = gate(px.bfloat16())
    assert w.shape == (6, 5) and t.shape == (6, 2)
    assert torch.allclose(w.float().sum(-1), torch.ones(6, device=DEV), atol=2e-2)
    scale = l32.abs().max().item()
    assert (l.float() - l32).abs().max().item() <= 5e-2 * scale


@pytest.mark.parametrize("size", [224, 96])
def test_native_resnet50_gate_vs_torchvision_fp32(size):
    """gating.B200ResNet50Gate: the reference gate's ResNet-50 with every convolution on the tcgen05 GEMM (row staging,
    BatchNorm folded, ReLU / add-then-ReLU epilogues, NHWC activations) against torchvision's own forward in fp32 — and
    against torchvision in bf16 on cuDNN (the reference's regime) for scale: the native path must not be further from
    fp32 than 2x what cuDNN's bf16 run is."""
    pytest.importorskip("torchvision")
    from multimeditron_b200.gating import B200ResNet50Gate, ResNet50Gate
    torch.manual_seed(5)
    ref = ResNet50Gate(5, top_k=2, device=DEV, dtype=torch.float32, fold_bn=False)
    for m in ref.resnet.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.normal_(0, 0.1); m.running_var.uniform_(0.5, 1.5)
            m.weight.data.uniform_(0.5, 1.5); m.bias.data.normal_(0, 0.1)
    sd = {"gating_network.resnet." + k: v for k, v in ref.resnet.state_dict().items()}
    gate = B200ResNet50Gate(sd, 5, top_k=2, device=DEV, prefix="gating_network.resnet.")
    lib16 = ResNet50Gate(5, top_k=2, device=DEV, state_dict=sd, prefix="gating_network.resnet.", fold_bn=False)
    px = torch.randn(6, 3, size, size, device=DEV)
    l32, t32, w32 = ref(px)
    l, t, w = gate(px)
    lb, _, _ = lib16(px.bfloat16())
    assert l.shape == (6, 5) and w.shape == (6, 5) and t.shape == (6, 2)
    scale = l32.abs().max().item()
    err, lib_err = (l.float() - l32).abs().max().item(), (lb.float() - l32).abs().max().item()
    print(f"[native gate {size}px] logits max-abs {err:.3e} (torchvision bf16 on cuDNN: {lib_err:.3e}) of scale {scale:.3e}; "
          f"weights max-abs {(w.float() - w32).abs().max().item():.3e}")
    assert err <= 2.0 * lib_err + 1e-2 * scale
    assert torch.allclose(w.float().sum(-1), torch.ones(6, device=DEV), atol=2e-2)
    lp, _, _ = gate(px.bfloat16())                    # bf16 pixels (what the device pre-processor hands over)
    assert (lp.float() - l32).abs().max().item() <= 2.0 * lib_err + 2e-2 * scale


def test_conv_rows_and_pools_bit_exact():
    """mm_conv_rows_nhwc / _stem, mm_maxpool_nhwc, mm_avgpool_nhwc against torch (unfold / max_pool2d / mean)."""
    from multimeditron_b200 import ops
    import torch.nn.functional as F
    g0 = torch.Generator().manual_seed(1)
    x = torch.randn(3, 10, 13, 16, generator=g0).bfloat16().to(DEV)                    # NHWC
    for (k, s, p) in ((3, 1, 1), (3, 2, 1), (1, 2, 0)):
        kpad = (k * k * 16 + 63) // 64 * 64
        rows = ops.conv_rows_nhwc(x, k, k, s, p, kpad)
        unf = F.unfold(x.permute(0, 3, 1, 2).float(), k, padding=p, stride=s)           # [B, C*k*k, L], (c, ky, kx) order
        L = unf.shape[-1]
        want = unf.view(3, 16, k * k, L).permute(0, 3, 2, 1).reshape(3 * L, k * k * 16)   # -> (ky,kx,c) order
        assert torch.equal(rows[:, :k * k * 16].float(), want) and not rows[:, k * k * 16:].any()
    px = torch.randn(2, 3, 20, 23, generator=g0).to(DEV)
    rows = ops.conv_rows_stem(px, 7, 7, 2, 3, 192)
    unf = F.unfold(px, 7, padding=3, stride=2)
    L = unf.shape[-1]
    want = unf.view(2, 3, 49, L).permute(0, 3, 2, 1).reshape(2 * L, 147).bfloat16()
    assert torch.equal(rows[:, :147], want) and not rows[:, 147:].any()
    mp = ops.maxpool_nhwc(x)
    assert torch.equal(mp.permute(0, 3, 1, 2).float(), F.max_pool2d(x.permute(0, 3, 1, 2).float(), 3, 2, 1))
    ap = ops.avgpool_nhwc(x.view(3, 130, 16))
    assert torch.equal(ap, x.view(3, 130, 16).float().mean(1).bfloat16())
    # ReLU epilogue applied AFTER the residual add
    a = torch.randn(300, 64, generator=g0).bfloat16().to(DEV)
    w = (torch.randn(64, 64, generator=g0) * 0.2).bfloat16().to(DEV)
    b = torch.randn(64, generator=g0).bfloat16().to(DEV)
    r = torch.randn(300, 64, generator=g0).bfloat16().to(DEV)
    got = ops.gemm_bias_act(a, w, b, act=ops.ACT_RELU, residual=r)
    lin = ops.gemm_bias_act(a, w, b)
    assert torch.equal(got, torch.relu((lin.float() + r.float()).bfloat16()))
    assert torch.equal(ops.gemm_bias_act(a, w, b, act=ops.ACT_RELU), torch.relu(lin))


def test_grouped_gemm_bit_identical_to_per_group_launches():
    """mm_gemm_grouped: G same-shaped GEMMs in one launch == G launches (plain, residual + stats producer, LN-fold consumer,
    row-mapped output), bit for bit; a row count that would let a tile straddle two groups is refused."""
    from multimeditron_b200 import ops
    G, M, N, K = 3, 512, 768, 256
    g0 = torch.Generator().manual_seed(9)
    a = torch.randn(G * M, K, generator=g0).bfloat16().to(DEV)
    w = (torch.randn(G * N, K, generator=g0) * 0.05).bfloat16().to(DEV)
    b = torch.randn(G * N, generator=g0).bfloat16().to(DEV)
    res = torch.randn(G * M, N, generator=g0).bfloat16().to(DEV)
    per = lambda fn: torch.cat([fn(g) for g in range(G)], 0)  # noqa: E731
    sl = lambda t, g, n: t[g * n:(g + 1) * n]                   # noqa: E731
    assert torch.equal(ops.gemm_grouped(a, w, b, G, act=1),
                       per(lambda g: ops.gemm_bias_act(sl(a, g, M), sl(w, g, N), sl(b, g, N), act=1)))
    st = torch.zeros(G * M, ops.row_stats_slots(N), 2, device=DEV)
    got = ops.gemm_grouped(a, w, b, G, residual=res, stats_out=st)
    st1 = torch.zeros_like(st)
    want = per(lambda g: ops.gemm_ln_bias_act(sl(a, g, M), sl(w, g, N), sl(b, g, N), residual=sl(res, g, M),
                                              stats_out=sl(st1, g, M)))
    assert torch.equal(got, want) and torch.equal(st, st1)
    lns = (torch.rand(G * M, ops.row_stats_slots(K), 2, generator=g0) * 64 + 0.5).to(DEV)
    assert torch.equal(ops.gemm_grouped(a, w, b, G, act=1, ln_stats=lns, eps=1e-5),
                       per(lambda g: ops.gemm_ln_bias_act(sl(a, g, M), sl(w, g, N), sl(b, g, N), act=1,
                                                          ln_stats=sl(lns, g, M), eps=1e-5)))
    rmap = torch.randperm(G * M, generator=g0).int().to(DEV)
    rmap[::7] = -1
    o1 = torch.full((G * M, N), 3.0, device=DEV).bfloat16()
    o2 = o1.clone()
    ops.gemm_grouped(a, w, b, G, out=o1, out_row_map=rmap)
    for g in range(G):
        ops.gemm_bias_act(sl(a, g, M), sl(w, g, N), sl(b, g, N), out=o2, out_row_map=sl(rmap, g, M))
    assert torch.equal(o1, o2)
    with pytest.raises(ops.MMError):
        ops.gemm_grouped(a[:G * 384].contiguous(), w, b, G)    # 384 rows per group: a 256-row tile would straddle groups


def test_grouped_experts_equal_per_expert_runners(monkeypatch):
    """vit.GroupedTowerRunner (one grouped launch per layer op for all experts) against the per-expert runners."""
    cfg = TOWERS["tiny"]        # T = 17: B*T % 256 == 0 needs B = 256
    E, hidden, B = 3, 64, 256
    names = [f"expert{i}" for i in range(E)]
    mcfg = AutoModality.config_from_dict({
        "model_type": "moe_meditron_clip_b200", "hidden_size": hidden, "expert_clip_names": names,
        "expert_towers": {n: "tiny" for n in names}, "image_processor": "tiny", "fusion_method": "weighted_average",
        "top_k_experts": 1, "gating_class_names": [names[i] for i in (2, 0, 1)]})
    sd = synth.make_moe_state_dict([cfg] * E, hidden, False, seed=4)
    w = torch.softmax(torch.randn(B, E, generator=torch.Generator().manual_seed(2)), -1)
    px = synth.make_pixels(cfg, B, seed=3).to(DEV)
    gate = lambda p: w[:p.shape[0]].to(p.device)     # noqa: E731
    m = AutoModality.from_config(mcfg).load_reference_state_dict(sd, DEV, gating=gate)
    assert m._grouped is not None and m._grouped.eligible(B)
    # the experts' own weights are row slices of the stacked tensors (one resident copy, not two)
    for i in (0, 1):
        for k in ("wo", "w2", "w1_f", "wqkv" if i == 0 else "wqkv_f"):
            stacked = m._grouped.layers[i][k]
            n = stacked.shape[0] // E
            for e in range(E):
                mine = m._runners[e].w.layers[i][k]
                assert mine.data_ptr() == stacked[e * n].data_ptr() and mine.shape[0] == n and mine.is_contiguous()
    y = m(px)
    monkeypatch.setenv("MM_B200_MOE_GROUPED", "0")
    m2 = AutoModality.from_config(mcfg).load_reference_state_dict(sd, DEV, gating=gate)
    assert m2._grouped is None
    assert torch.equal(y, m2(px))
    assert not m._grouped.eligible(3) and torch.equal(m(px[:3]), m2(px[:3]))     # ineligible batch: per-expert path
