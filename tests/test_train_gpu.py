"""-m gpu: the projector's training step on the kernels and the small kernels around it.

Reference: MLPProjector (model/projectors/mlp.py:33-39) under torch.autograd, as the ALIGNMENT stage runs it
(model/model.py:310-326).  Ground truth = autograd through oracle.vit_oracle.mlp_projector in fp32 on the GPU; the
reference's own bf16 regime gives the error scale.  Tolerance: |g - g32| <= 2 x |g_bf16 - g32| + 1e-2 * max|g32|.
"""
import pytest
import torch
import torch.nn.functional as F

from multimeditron_b200 import autograd_ops, ops, synth
from multimeditron_b200.configs import TOWERS
from multimeditron_b200.modalities import AutoModality
from multimeditron_b200.model import B200MultiModalEmbedder
from oracle import vit_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(got, ref, tol=2e-2, ulp=1.6e-2):
    got, ref = got.float(), ref.float()
    bound = tol * ref.abs().max() + ulp * ref.abs()
    bad = (got - ref).abs() > bound
    assert not bad.any(), f"{int(bad.sum())} elements off, max err {(got - ref).abs().max():.3e}, scale {ref.abs().max():.3e}"


@pytest.mark.parametrize("R,C", [(64, 64), (72, 200), (1000, 1024), (4096, 1152), (147_456 // 8, 4096)])
def test_transpose_bit_exact(R, C):
    x = torch.randn(R, C, device=DEV).bfloat16()
    assert torch.equal(ops.transpose(x), x.t().contiguous())


@pytest.mark.parametrize("n", [8, 4096, 577 * 4096])
def test_gelu_fwd_matches_fused_epilogue_and_torch(n):
    z = (torch.randn(n, device=DEV) * 3).bfloat16()
    a = ops.gelu_fwd(z)
    _close(a, F.gelu(z.float()), tol=1e-3, ulp=8e-3)
    if n % 256 == 0:   # bit-identical to the GEMM's fused erf-GELU epilogue: identity weight, zero bias
        K = 256
        zz = z.view(-1, K)[:300].contiguous()
        eye = torch.eye(K, device=DEV).bfloat16()
        fused = ops.gemm_bias_act(zz, eye, None, act=ops.ACT_GELU_ERF)
        assert torch.equal(fused, ops.gelu_fwd(zz))


@pytest.mark.parametrize("M,N", [(8, 8), (301, 264), (5000, 1024), (4617, 4096)])
def test_gelu_bwd_and_column_sums(M, N):
    g = torch.Generator(device=DEV).manual_seed(M + N)
    da = torch.randn(M, N, device=DEV, generator=g).bfloat16()
    z = (torch.randn(M, N, device=DEV, generator=g) * 2).bfloat16()
    dz, db = ops.gelu_bwd_colsum(da, z)
    zf = z.float().requires_grad_(True)
    F.gelu(zf).backward(da.float())
    _close(dz, zf.grad, tol=1e-3, ulp=8e-3)
    want = dz.float().sum(0)                       # the sum runs over the stored bf16 values
    assert torch.allclose(db, want, rtol=1e-4, atol=1e-3 * max(1.0, M ** 0.5))
    dz2, db2 = ops.gelu_bwd_colsum(da, z)
    assert torch.equal(db, db2) and torch.equal(dz, dz2)     # deterministic
    none, cs = ops.gelu_bwd_colsum(da, None)
    assert none is None and torch.allclose(cs, da.float().sum(0), rtol=1e-4, atol=1e-3 * max(1.0, M ** 0.5))


def test_gather_rows_is_the_scatter_backward():
    H, n_dest, n = 512, 4000, 1500
    src = torch.randn(n_dest, H, device=DEV).bfloat16()
    rows = torch.randperm(n_dest, device=DEV)[:n].to(torch.int32)
    assert torch.equal(ops.gather_rows(src, rows), src[rows.long()])
    assert torch.equal(ops.gather_rows(src, rows, out_dtype=torch.float32), src[rows.long()].float())
    st = ops.new_status(DEV)
    rows[3] = n_dest
    ops.gather_rows(src, rows, status=st)
    with pytest.raises(IndexError):
        ops.raise_for_status(st)


def test_chained_splices_keep_every_source_gradient():
    """Two splices into one sequence (two modality types): the second one's base is the first one's output, so its
    backward must hand the base a gradient (zero on the rows it overwrote) — checked against index_put_ autograd,
    including a row both splices write (the later one wins, the earlier source gets zero there)."""
    H, n_dest = 256, 600
    g = torch.Generator(device=DEV).manual_seed(11)
    base = torch.randn(n_dest, H, device=DEV, generator=g).bfloat16()
    perm = torch.randperm(n_dest, device=DEV, generator=g)
    r1, r2 = perm[:150].clone(), perm[150:260].clone()
    r2[0] = r1[5]                                                     # overlap
    s1 = torch.randn(150, H, device=DEV, generator=g).bfloat16()
    s2 = torch.randn(110, H, device=DEV, generator=g).bfloat16()
    w = torch.randn(n_dest, H, device=DEV, generator=g).bfloat16()

    a1, a2 = s1.clone().requires_grad_(True), s2.clone().requires_grad_(True)
    out = autograd_ops.splice_rows(base.clone(), a1, r1)
    out = autograd_ops.splice_rows(out.view(2, n_dest // 2, H).view(n_dest, H), a2, r2)   # through views, like model.py
    (out.float() * w.float()).sum().backward()

    b1, b2 = s1.clone().requires_grad_(True), s2.clone().requires_grad_(True)
    ref = base.clone()
    ref = ref.index_put((r1,), b1)
    ref = ref.index_put((r2,), b2)
    (ref.float() * w.float()).sum().backward()
    assert torch.equal(out.detach(), ref.detach())
    assert a1.grad is not None and torch.equal(a1.grad, b1.grad)
    assert torch.equal(a2.grad, b2.grad)
    assert not a1.grad[5].any() and a1.grad[4].any()


def test_two_trainable_modalities_alignment_step():
    """ALIGNMENT stage with two modality types whose projectors both train (collator positions): each projector's
    gradients against fp32 autograd through the oracle projector.  Before the fix the first-spliced modality's
    projector silently received no gradient."""
    cfg = TOWERS["tiny"]
    H, V, B, S = 64, 300, 2, 64
    att, P = V - 4, cfg.num_patches
    mods, sds = [], []
    for i, name in enumerate(("image", "scan")):
        sd = synth.make_modality_state_dict(cfg, H, seed=20 + i)
        m = AutoModality.from_config(AutoModality.config_from_dict(
            {"model_type": "meditron_clip_b200", "hidden_size": H, "clip_name": "tiny"})).load_reference_state_dict(sd, DEV)
        m.config.modality_type = name
        mods.append(m)
        sds.append(sd)
    # per sequence: [text, P attachment tokens of "image", text, P attachment tokens of "scan", text]
    assert 2 * P + 6 <= S
    ids = torch.randint(0, V - 8, (B, S))
    o1, o2 = 2, 2 + P + 2
    ids[:, o1:o1 + P] = att
    ids[:, o2:o2 + P] = att
    px = [synth.make_pixels(cfg, B, seed=30 + i) for i in range(2)]
    bidx = torch.arange(B).repeat_interleave(P)
    pmi = {"stacked": {"image": [px[0][b] for b in range(B)], "scan": [px[1][b] for b in range(B)]},
           "batch_idx": {"image": bidx, "scan": bidx.clone()},
           "token_range": {"image": (o1 + torch.arange(P)).repeat(B), "scan": (o2 + torch.arange(P)).repeat(B)}}
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(3)) * 0.02).bfloat16().to(DEV)
    emb = B200MultiModalEmbedder(table, mods, att)
    with torch.no_grad():
        base = emb(ids, pmi, positions_from="collator").clone()
    for m in mods:
        m.unfreeze_projection()
    out = emb(ids, pmi, positions_from="collator")
    emb.check()
    assert out.requires_grad and torch.equal(out.detach(), base)
    w = torch.randn_like(out)
    (out.float() * w.float()).sum().backward()
    for i, (m, off) in enumerate(zip(mods, (o1, o2))):
        with torch.no_grad():
            feats = m.runner.features(px[i].to(DEV)).float()
        p32 = {k: v.to(DEV).float().requires_grad_(True) for k, v in sds[i].items() if k.startswith("projector.")}
        y = vit_oracle.mlp_projector(p32, feats)
        y.backward(w[:, off:off + P].reshape(-1, H).float())
        for k, p in m.named_parameters():
            assert p.grad is not None, (m.config.modality_type, k)
            r = p32[k].grad
            err, scale = (p.grad.float() - r).abs().max().item(), r.abs().max().item()
            assert err <= 3e-2 * scale + 1e-6, (m.config.modality_type, k, err, scale)


def test_embed_gather_skip_leaves_attachment_rows_alone():
    V, H, att = 300, 256, 296
    table = torch.randn(V, H, device=DEV).bfloat16()
    ids = torch.randint(0, V - 8, (4, 50), device=DEV)
    ids[:, 10:30] = att
    full = ops.embed_gather(table, ids)
    part = ops.embed_gather(table, ids, skip_id=att)
    keep = ids != att
    assert torch.equal(part[keep], full[keep]) and torch.equal(full[~keep], table[att].expand(int((~keep).sum()), H))


@pytest.mark.parametrize("N,K", [(256, 128), (3072, 1024), (4304, 1152), (1152, 4304)])
def test_fold_layernorm_kernel(N, K):
    g0 = torch.Generator().manual_seed(N + K)
    w = (torch.randn(N, K, generator=g0) * 0.02).bfloat16().to(DEV)
    b = (torch.randn(N, generator=g0) * 0.02).bfloat16().to(DEV)
    gamma = (1 + 0.1 * torch.randn(K, generator=g0)).bfloat16().to(DEV)
    beta = (0.1 * torch.randn(K, generator=g0)).bfloat16().to(DEV)
    q, bo = ops.fold_layernorm(w, b, gamma, beta)
    wg = w.double() * gamma.double()[None]
    wg = wg - wg.mean(1, keepdim=True)
    import math
    ulp = 2.0 ** (math.floor(math.log2(wg.abs().max().item())) - 7)      # bf16 spacing in the binade of the largest |w|
    # sigma-delta rounding: |q - wg| <= |q - t| + |carried error| <= ulp/2 + ulp/2
    assert (q.double() - wg).abs().max().item() <= 1.01 * ulp
    assert q.double().sum(1).abs().max().item() <= 0.55 * ulp + 1e-6     # ... and the ROW SUMS stay within half an ulp of zero
    want_b = (w.double() @ beta.double() + b.double())
    assert (bo.double() - want_b).abs().max().item() <= want_b.abs().max().item() * 2 ** -8


@pytest.mark.parametrize("M,D,H", [(98, 128, 64), (8 * 49, 768, 512), (4 * 576, 1024, 4096)])
def test_projector_forward_backward_vs_fp32_autograd(M, D, H):
    sd = {k: v.to(DEV) for k, v in synth._draw(synth.projector_keys(D, H), torch.Generator().manual_seed(M), torch.float32, "cpu").items()}
    x = torch.randn(M, D, device=DEV) * 0.7
    dy = torch.randn(M, H, device=DEV) * 0.1
    names = [f"projector.projection.{i}.{n}" for i in (0, 2, 4) for n in ("weight", "bias")]

    def run(dtype):
        p = {k: sd[k].detach().clone().to(dtype).requires_grad_(True) for k in names}     # fresh leaves per dtype
        xx = x.detach().clone().to(dtype).requires_grad_(True)
        y = vit_oracle.mlp_projector(p, xx)
        y.backward(dy.to(dtype))
        return y.detach().float(), [xx.grad.float()] + [p[k].grad.float() for k in names]

    y32, g32 = run(torch.float32)
    ybf, gbf = run(torch.bfloat16)
    params = [sd[k].bfloat16().requires_grad_(True) for k in names]
    xb = x.bfloat16().requires_grad_(True)
    y = autograd_ops.ProjectorFunction.apply(xb, *params)
    # forward: identical rounding points to the inference path (fused GELU epilogues)
    a = ops.gemm_bias_act(xb.detach(), params[0].detach(), params[1].detach(), act=ops.ACT_GELU_ERF)
    a = ops.gemm_bias_act(a, params[2].detach(), params[3].detach(), act=ops.ACT_GELU_ERF)
    assert torch.equal(y.detach(), ops.gemm_bias_act(a, params[4].detach(), params[5].detach()))
    y.backward(dy.bfloat16())
    got = [xb.grad.float()] + [p.grad.float() for p in params]
    _close(y.detach(), y32)
    for name, g, r32, rbf in zip(["x"] + names, got, g32, gbf):
        scale = r32.abs().max().item()
        err, ref_err = (g - r32).abs().max().item(), (rbf - r32).abs().max().item()
        print(f"[projector bwd M={M} D={D} H={H}] {name}: max-abs {err:.3e} (torch bf16 autograd: {ref_err:.3e}) /absmax {err / scale:.3e}")
        assert err <= 2.0 * ref_err + 1e-2 * scale, name


def test_embedder_alignment_step_mirror_class():
    """The repo's own mirror (no reference installed): unfreeze_projection() exposes nn.Parameters that share storage
    with the kernel-side weights; one forward+backward through embed_modalities_with_text, then an SGD step that the
    next (inference) forward must see."""
    cfg = TOWERS["tiny"]
    H, V, B, S = 64, 300, 2, 48
    sd = synth.make_modality_state_dict(cfg, H, seed=2)
    m = AutoModality.from_config(AutoModality.config_from_dict(
        {"model_type": "meditron_clip_b200", "hidden_size": H, "clip_name": "tiny"})).load_reference_state_dict(sd, DEV)
    att = V - 4
    ids, _ = synth.make_sequences(B, S, 1, cfg.num_patches, V - 4, att, V - 3, V - 2, V - 1, seed=1)
    px = synth.make_pixels(cfg, B, seed=1)
    pmi = {"stacked": {"image": [px[i] for i in range(B)]}, "batch_idx": {}, "token_range": {}}
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(3)) * 0.02).bfloat16().to(DEV)
    emb = B200MultiModalEmbedder(table, [m], att)
    with torch.no_grad():
        base = emb(ids, pmi, positions_from="device").clone()
    m.unfreeze_projection()
    assert sorted(k for k, _ in m.named_parameters()) == sorted(
        f"projector.projection.{i}.{n}" for i in (0, 2, 4) for n in ("weight", "bias"))
    out = emb(ids, pmi, positions_from="device")
    emb.check()
    assert out.requires_grad and torch.equal(out.detach(), base)          # same bits as the fused inference path
    w = torch.randn_like(out)
    (out.float() * w.float()).sum().backward()
    # reference gradient: fp32 autograd through the oracle projector on the same features
    with torch.no_grad():
        feats = m.runner.features(px.to(DEV)).float()
    p32 = {k: v.to(DEV).float().requires_grad_(True) for k, v in sd.items() if k.startswith("projector.")}
    y = vit_oracle.mlp_projector(p32, feats)
    rows = (ids.to(DEV).view(-1) == att).nonzero().squeeze(1)
    y.backward(w.view(-1, H)[rows].float())
    for k, p in m.named_parameters():
        r = p32[k].grad
        err, scale = (p.grad.float() - r).abs().max().item(), r.abs().max().item()
        assert err <= 3e-2 * scale + 1e-6, (k, err, scale)
    with torch.no_grad():
        for p in m.parameters():
            p -= 0.5 * p.grad
    m.freeze_projection()
    with torch.no_grad():
        after = emb(ids, pmi, positions_from="device")
    assert not torch.equal(after, base)                                    # the kernels multiply by the updated weights


@pytest.mark.parametrize("pep,fusion", [(False, "weighted_average"), (False, "sequence_append"),
                                        (True, "weighted_average"), (True, "sequence_append")])
def test_moe_alignment_stage_projector_gradients(pep, fusion):
    """ALIGNMENT stage of the MoE modalities (freeze_for_alignment -> unfreeze_projection + freeze_modality_embedder,
    model.py:310-326; image_modality_moe.py:222-246, image_modality_moe_pep.py:263-288): experts / gate / fusion on the
    kernels without a graph, the projector(s) under autograd.  Gradients of every projector tensor against torch
    autograd through the oracle (fp32 ground truth, bf16 = the reference's own regime)."""
    from multimeditron_b200.configs import TOWERS
    from multimeditron_b200.modalities import AutoModality
    cfg, E, hidden, B = TOWERS["tiny"], 3, 64, 5
    names = [f"expert{i}" for i in range(E)]
    perm = [2, 0, 1]
    mcfg = AutoModality.config_from_dict({
        "model_type": "moe_meditron_clip_pep_b200" if pep else "moe_meditron_clip_b200", "hidden_size": hidden,
        "expert_clip_names": names, "expert_towers": {n: "tiny" for n in names}, "image_processor": "tiny",
        "fusion_method": fusion, "top_k_experts": E if fusion == "sequence_append" else 1,
        "gating_class_names": [names[i] for i in perm]})
    sd = synth.make_moe_state_dict([cfg] * E, hidden, pep, seed=12)
    w = torch.softmax(torch.randn(B, E, generator=torch.Generator().manual_seed(1)), -1)
    px = synth.make_pixels(cfg, B, seed=2)
    m = AutoModality.from_config(mcfg).load_reference_state_dict(sd, DEV, gating=lambda p: w.to(p.device))
    m.unfreeze_projection()
    m.freeze_modality_embedder()
    assert m.projector_trainable
    y = m([px[i] for i in range(B)])
    assert y.requires_grad
    dy = (torch.randn(y.shape, generator=torch.Generator().manual_seed(3)) * 0.1).to(DEV)
    y.backward(dy.to(y.dtype))
    pkeys = [k for k in sd if k.startswith("projector.") or k.startswith("projectors.")]
    mods = m.projectors if pep else m.projector
    got = {("projectors." if pep else "projector.") + n: p.grad.float() for n, p in mods.named_parameters()}
    assert set(got) == set(pkeys), (sorted(got)[:3], sorted(pkeys)[:3])

    def run(dtype):
        p = {k: (v.detach().clone().to(DEV, dtype).requires_grad_(True) if k in pkeys else v.to(DEV, dtype))
             for k, v in sd.items()}
        out = vit_oracle.moe_modality_forward(p, [cfg] * E, px.to(DEV, dtype), w.to(DEV, dtype), perm, fusion, pep)
        out.backward(dy.to(dtype))
        return out.detach().float(), {k: p[k].grad.float() for k in pkeys}

    y32, g32 = run(torch.float32)
    ybf, gbf = run(torch.bfloat16)
    _close(y.detach(), y32)
    for k in pkeys:
        scale = g32[k].abs().max().item()
        err, ref_err = (got[k] - g32[k]).abs().max().item(), (gbf[k] - g32[k]).abs().max().item()
        assert err <= 2.0 * ref_err + 2e-2 * scale, (k, err, ref_err, scale)
    # one optimizer step moves the weights the NEXT forward multiplies by (shared storage)
    opt = torch.optim.SGD(mods.parameters(), lr=0.5)
    opt.step()
    with torch.no_grad():
        y2 = m([px[i] for i in range(B)])
    assert not torch.equal(y2, y.detach())
    with pytest.raises(NotImplementedError):
        m.unfreeze_modality_embedder()
