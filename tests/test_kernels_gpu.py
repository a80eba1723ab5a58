"""-m gpu: floating-point kernels through the C-ABI vs a plain PyTorch fp32 reference of the same op.

Tolerance (stated, bf16 storage / fp32 accumulate): |got-ref| <= 2e-2*max|ref| + 1.6e-2*|ref|  (2 bf16 ulps of the
largest value + 2 ulps relative) — one bf16 rounding of the output plus bf16-rounded intermediates."""
import pytest
import torch
import torch.nn.functional as F

from multimeditron_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def close(got, ref, tol=2e-2):
    got, ref = got.float(), ref.float()
    assert torch.isfinite(got).all()
    err = (got - ref).abs()
    lim = tol * ref.abs().max() + 1.6e-2 * ref.abs()
    bad = err > lim
    assert not bad.any(), f"max err {err.max().item():.3e}, {int(bad.sum())} bad of {bad.numel()}"


@pytest.mark.parametrize("M,D", [(1, 128), (77, 768), (300, 1024), (50, 1152), (4096, 1024)])
def test_layernorm(M, D):
    g0 = torch.Generator().manual_seed(M * D)
    x = (torch.randn(M, D, generator=g0) * 3 + 1).bfloat16().to(DEV)
    g = (1 + 0.1 * torch.randn(D, generator=g0)).bfloat16().to(DEV)
    b = (0.1 * torch.randn(D, generator=g0)).bfloat16().to(DEV)
    for eps in (1e-5, 1e-6):
        y = ops.layernorm(x, g, b, eps)
        close(y, F.layer_norm(x.float(), (D,), g.float(), b.float(), eps))


def test_layernorm_row_remap_drops_cls():
    B, T, D = 3, 17, 128
    x = torch.randn(B * T, D, device=DEV).bfloat16()
    g = torch.ones(D, device=DEV).bfloat16()
    b = torch.zeros(D, device=DEV).bfloat16()
    out = torch.empty(B * (T - 1), D, device=DEV).bfloat16()
    ops.layernorm(x, g, b, 1e-5, out=out, rows_in_group=T, rows_out_group=T - 1, in_row_offset=1)
    ref = F.layer_norm(x.float().view(B, T, D)[:, 1:], (D,)).reshape(-1, D)
    close(out, ref)


GEMM_CASES = [
    # M, N, K, bias, act, residual, rowmap
    (128, 256, 64, False, 0, False, False),
    (128, 256, 1024, False, 0, False, False),      # ring wrap
    (300, 264, 200, False, 0, False, False),       # M/N/K tails
    (1000, 768, 3072, True, 1, False, False),      # CLIP fc1 + QuickGELU
    (1000, 1024, 1024, True, 0, True, False),      # out_proj + residual
    (1731, 1024, 4096, True, 2, True, True),       # erf-GELU + residual + row map
    (50, 1152, 4304, True, 3, False, False),       # SigLIP shapes, tanh-GELU
    (400, 2304, 768, True, 0, False, False),       # cfg1 fused QKV
    (56_837, 512, 128, True, 0, False, False),     # many tiles per CTA (TMEM double buffer, persistence)
    (8, 8, 8, True, 0, False, False),              # minimum
]


@pytest.mark.parametrize("M,N,K,bias,act,residual,rowmap", GEMM_CASES)
def test_gemm_bias_act(M, N, K, bias, act, residual, rowmap):
    g0 = torch.Generator().manual_seed(M + 7 * N + 13 * K)
    a = torch.randn(M, K, generator=g0).bfloat16().to(DEV)
    w = (torch.randn(N, K, generator=g0) * 0.05).bfloat16().to(DEV)
    bs = (torch.randn(N, generator=g0) * 0.5).bfloat16().to(DEV) if bias else None
    res = torch.randn(M, N, generator=g0).bfloat16().to(DEV) if residual else None
    ref = a.float() @ w.float().t()
    if bias:
        ref = ref + bs.float()
    ref = {0: lambda t: t, 1: lambda t: t * torch.sigmoid(1.702 * t), 2: F.gelu,
           3: lambda t: F.gelu(t, approximate="tanh")}[act](ref)
    if residual:
        ref = ref + res.float()
    if rowmap:
        perm = torch.randperm(M, generator=g0).int().to(DEV)
        perm[::7] = -1
        out = torch.full((M, N), 7.0, device=DEV).bfloat16()
        ops.gemm_bias_act(a, w, bs, act, residual=res, out=out, out_row_map=perm)
        want = torch.full((M, N), 7.0, device=DEV)
        keep = perm >= 0
        want[perm[keep].long()] = ref[keep]
        close(out, want)
    else:
        close(ops.gemm_bias_act(a, w, bs, act, residual=res), ref)


@pytest.mark.parametrize("M,N,K,bias,residual", [
    (1000, 1024, 1024, True, False),      # out-proj shape, pair kernel
    (577, 4096, 1024, True, False),       # fc1 shape (no activation: the bound below is about the accumulation)
    (300, 1024, 4096, False, True),       # fc2 shape: long k, residual add
    (128, 1152, 4304, True, False),       # M <= 128 -> 1-CTA kernel, k not a multiple of 64
])
def test_gemm_elementwise_error_bound(M, N, K, bias, residual):
    """Per-ELEMENT bound, much tighter than `close` (which allows ~5 ulps of the LARGEST output on every element, so a
    dropped k-slab on a small-magnitude row could slip through): against a float64 product of the same bf16 operands the
    stored value may differ by the final bf16 rounding(s) (1 ulp of the element = 2^-8 |ref|, twice with a residual) plus
    the fp32 accumulation error, which is bounded by a small multiple of sum_k |a_k w_k| — 2e-4 here, while a lost
    64-wide k-slab would move an element by ~64/K of that sum (>= 1.5e-2)."""
    g0 = torch.Generator().manual_seed(3 * M + 5 * N + 7 * K)
    a = torch.randn(M, K, generator=g0).bfloat16().to(DEV)
    a[::3] *= 0.01                                           # small-magnitude rows: invisible to a max-scaled tolerance
    w = (torch.randn(N, K, generator=g0) * 0.05).bfloat16().to(DEV)
    bs = (torch.randn(N, generator=g0) * 0.02).bfloat16().to(DEV) if bias else None
    res = (torch.randn(M, N, generator=g0) * 0.05).bfloat16().to(DEV) if residual else None
    got = ops.gemm_bias_act(a, w, bs, residual=res).double()
    ref = a.double() @ w.double().t()
    mag = a.double().abs() @ w.double().abs().t()            # sum_k |a_k w_k|
    if bias:
        ref = ref + bs.double()
        mag = mag + bs.double().abs()
    lim = 2.0 ** -8 * ref.abs() + 2e-4 * mag
    if residual:                                             # bf16(bf16(acc + bias) + res): a second rounding
        lim = lim + 2.0 ** -8 * (ref + res.double()).abs()
        ref = ref + res.double()
    err = (got - ref).abs()
    bad = err > lim
    assert not bad.any(), f"{int(bad.sum())} of {bad.numel()} elements outside the bound, worst ratio {(err / lim).max().item():.2f}"
    # the bound does catch a dropped k-slab (the check checks something): zero 64 columns of k and compare again
    a2 = a.clone()
    a2[:, 64:128] = 0
    got2 = ops.gemm_bias_act(a2, w, bs, residual=res).double()
    assert ((got2 - ref).abs() > lim).float().mean().item() > 0.5


def test_gemm_modular_residual_and_strided_output():
    M, N, K, P = 96, 128, 64, 16
    a = torch.randn(M, K, device=DEV).bfloat16()
    w = (torch.randn(N, K, device=DEV) * 0.1).bfloat16()
    pos = torch.randn(P + 1, N, device=DEV).bfloat16()
    big = torch.zeros(M, 2 * N, device=DEV).bfloat16()
    ops.gemm_bias_act(a, w, residual=pos, res_mod=P, res_row_off=1, out=big[:, N:])
    m = torch.arange(M, device=DEV)
    close(big[:, N:], a.float() @ w.float().t() + pos.float()[(m % P) + 1])
    assert not big[:, :N].any()


@pytest.mark.parametrize("M,D,N,act,eps,shift", [
    (300, 128, 256, 1, 1e-5, 0.0),        # tiny tower width, one stats slot pair
    (1000, 1024, 3072, 0, 1e-5, 0.0),     # CLIP-L QKV
    (577, 1024, 4096, 1, 1e-5, 3.0),      # CLIP-L fc1 + QuickGELU, rows with a large mean (cancellation in acc - mean*c)
    (729, 1152, 4304, 3, 1e-6, 0.5),      # SigLIP fc1 + tanh-GELU, 18 slots
    (130, 144, 480, 0, 1e-6, 0.0),        # widths that are not multiples of 64/128: partial last chunk
    (600, 1024, 1024, 1, 1e-5, -40.0),    # "massive activation" rows: |mean| >> spread of most channels (E[x^2]-mean^2 in fp32)
])
def test_gemm_layernorm_fold(M, D, N, act, eps, shift):
    """Producer GEMM (+bias +residual) emits per-chunk row statistics; consumer GEMM applies LayerNorm from them.
    Compared with fp32 torch: LN(x) @ W.T + b (x = what the producer stored)."""
    g0 = torch.Generator().manual_seed(M + N)
    a = torch.randn(M, 256, generator=g0).bfloat16().to(DEV)
    wo = (torch.randn(D, 256, generator=g0) * 0.05).bfloat16().to(DEV)
    bo = (torch.randn(D, generator=g0) * 0.1).bfloat16().to(DEV)
    outlier = 200.0 if abs(shift) > 10 else 4.0      # a few channels carry huge values, as in trained CLIP towers
    res = (torch.randn(M, D, generator=g0) * (1.0 + outlier * (torch.rand(1, D, generator=g0) > 0.97)) + shift).bfloat16().to(DEV)
    stats = torch.zeros(M, ops.row_stats_slots(D), 2, device=DEV)   # contract: zero-initialised once
    assert ops.row_stats_slots(D) == 16
    x = ops.gemm_ln_bias_act(a, wo, bo, residual=res, stats_out=stats)
    x_plain = ops.gemm_bias_act(a, wo, bo, residual=res)
    assert torch.equal(x, x_plain)                               # emitting statistics must not change the output
    xs = x.float()
    # statistics: per-128-column sums of the stored bf16 values
    pad = (-D) % 128
    xp = torch.nn.functional.pad(xs, (0, pad)).view(M, -1, 128)       # one slot per 128 output columns
    nc = xp.shape[1]
    assert torch.allclose(stats[:, :nc, 0], xp.sum(-1), rtol=1e-5, atol=1e-3)
    assert torch.allclose(stats[:, :nc, 1], (xp * xp).sum(-1), rtol=1e-5, atol=1e-3)
    assert (stats[:, nc:] == 0).all()
    gamma = (1 + 0.1 * torch.randn(D, generator=g0)).bfloat16().to(DEV)
    beta = (0.1 * torch.randn(D, generator=g0)).bfloat16().to(DEV)
    w = (torch.randn(N, D, generator=g0) * 0.02).bfloat16().to(DEV)
    b = (torch.randn(N, generator=g0) * 0.02).bfloat16().to(DEV)
    wf, bf = ops.fold_layernorm(w, b, gamma, beta)
    y = ops.gemm_ln_bias_act(x, wf, bf, act=act, ln_stats=stats, eps=eps)
    h = F.layer_norm(xs, (D,), gamma.float(), beta.float(), eps)
    ref = h @ w.float().T + b.float()
    ref = {0: lambda t: t, 1: lambda t: t * torch.sigmoid(1.702 * t), 3: lambda t: F.gelu(t, approximate="tanh")}[act](ref)
    close(y, ref)
    # and it must agree with the unfolded two-kernel path at bf16 level
    y2 = ops.gemm_bias_act(ops.layernorm(x, gamma, beta, eps), w, b, act=act)
    close(y, y2.float())
    y_again = ops.gemm_ln_bias_act(x, wf, bf, act=act, ln_stats=stats, eps=eps)
    assert torch.equal(y, y_again)                               # fixed summation order: bit-reproducible


@pytest.mark.parametrize("M,N,K", [(129, 256, 64), (300, 264, 200), (1000, 1024, 1024), (2308, 3072, 1024),
                                   (56_837, 512, 128), (4617, 4096, 1024)])
def test_gemm_pair_kernel_bit_identical_to_single_cta(M, N, K):
    """The CTA-pair kernel (cta_group::2, 256 x 256 tiles; default for M > 128) against the 1-CTA kernel: every epilogue
    variant — plain, activation, residual + row map, statistics producer, LayerNorm-fold consumer — bit for bit."""
    g0 = torch.Generator().manual_seed(M ^ N ^ K)
    a = torch.randn(M, K, generator=g0).bfloat16().to(DEV)
    w = (torch.randn(N, K, generator=g0) * 0.05).bfloat16().to(DEV)
    bs = (torch.randn(N, generator=g0) * 0.5).bfloat16().to(DEV)
    res = torch.randn(M, N, generator=g0).bfloat16().to(DEV)
    perm = torch.randperm(M, generator=g0).int().to(DEV)
    perm[::5] = -1
    gamma, beta = (1 + 0.1 * torch.randn(K, generator=g0)).bfloat16().to(DEV), (0.1 * torch.randn(K, generator=g0)).bfloat16().to(DEV)
    wf, bf = ops.fold_layernorm(w, bs, gamma, beta)
    ln_stats = (torch.rand(M, 16, 2, generator=g0) * 64 + 0.5).to(DEV)
    got = {}
    for pair in (0, 1):
        f = 0 if pair else ops.GEMM_FLAG_1CTA      # per-call flag in `act` (no process-wide switch)
        o_map = torch.full((M, N), 7.0, device=DEV).bfloat16()
        ops.gemm_bias_act(a, w, bs, 2 | f, residual=res, out=o_map, out_row_map=perm)
        outs = [ops.gemm_bias_act(a, w, None, f), ops.gemm_bias_act(a, w, bs, 1 | f), o_map]
        if N <= 2048:
            st = torch.zeros(M, 16, 2, device=DEV)
            outs += [ops.gemm_ln_bias_act(a, w, bs, act=f, residual=res, stats_out=st), st]
        if K <= 2048:
            outs.append(ops.gemm_ln_bias_act(a, wf, bf, act=1 | f, ln_stats=ln_stats, eps=1e-5))
        got[pair] = outs
    for x, y in zip(got[0], got[1]):
        assert torch.equal(x, y)
    close(got[1][0], a.float() @ w.float().t())


def test_gemm_rejects_bad_arguments():
    a = torch.zeros(16, 12, device=DEV).bfloat16()   # K % 8 != 0
    w = torch.zeros(16, 12, device=DEV).bfloat16()
    with pytest.raises(ops.MMError):
        ops.gemm_bias_act(a, w)


@pytest.mark.parametrize("B,img,p,D,cls,bias,dt", [
    (2, 32, 8, 128, True, False, torch.float32),
    (3, 224, 32, 768, True, False, torch.float32),     # CLIP-B/32
    (2, 336, 14, 1024, True, False, torch.float32),    # CLIP-L/14-336 (K=588, padded to 640)
    (2, 56, 14, 256, False, True, torch.bfloat16),     # SigLIP-style: conv bias, no CLS, bf16 pixels
    (2, 384, 14, 1152, False, True, torch.float32),    # SigLIP so400m: 384 % 14 != 0, the conv drops the last 6 pixels
    (3, 61, 14, 128, True, False, torch.bfloat16),     # odd image size, remainder 5
    (9, 336, 14, 1024, True, False, torch.float32),    # many row tiles: persistent pairs, modular position rows
])
def test_patch_embed(B, img, p, D, cls, bias, dt):
    G = img // p
    T = G * G + (1 if cls else 0)
    g0 = torch.Generator().manual_seed(img * p)
    px = torch.randn(B, 3, img, img, generator=g0).to(dt).to(DEV)
    w = (torch.randn(D, 3, p, p, generator=g0) * 0.02).bfloat16().to(DEV)
    bs = (torch.randn(D, generator=g0) * 0.1).bfloat16().to(DEV) if bias else None
    pos = (torch.randn(T, D, generator=g0) * 0.1).bfloat16().to(DEV)
    c = (torch.randn(D, generator=g0) * 0.1).bfloat16().to(DEV) if cls else None
    K = 3 * p * p
    wp = torch.zeros(D, (K + 63) // 64 * 64, device=DEV).bfloat16()
    wp[:, :K] = w.reshape(D, -1)
    out = ops.patch_embed(px, wp, bs, pos, c, p)                       # staged patch rows + CTA-pair GEMM
    out_gather = ops.patch_embed(px, wp, bs, pos, c, p, staged=False)  # in-kernel gather (mm_patch_embed)
    assert torch.equal(out, out_gather)                                # same operands, same k order: bit-identical
    ref = F.conv2d(px.bfloat16().float(), w.float(), bs.float() if bias else None, stride=p).flatten(2).transpose(1, 2)
    if cls:
        ref = torch.cat([c.float().expand(B, 1, D), ref], 1)
    close(out, ref + pos.float())


@pytest.mark.parametrize("B,T,nh,hd", [
    (1, 1, 1, 64), (1, 50, 2, 64), (2, 128, 2, 64), (2, 129, 2, 64), (2, 200, 3, 64), (2, 577, 4, 64), (3, 640, 2, 64),
    (40, 577, 16, 64), (2, 769, 2, 64), (1, 1024, 3, 64),                       # > 6 kv tiles: streamed K/V ring
    (1, 1, 1, 72), (2, 96, 2, 72), (2, 97, 2, 72), (3, 300, 3, 72), (2, 729, 16, 72), (20, 729, 16, 72),  # SigLIP
    (1, 1000, 2, 72)])
def test_attention_fwd(B, T, nh, hd):
    hs = ops.attn_head_stride(hd)
    g0 = torch.Generator().manual_seed(B * T * nh)
    qkv = torch.zeros(B * T, 3, nh, hs, dtype=torch.bfloat16)
    qkv[..., :hd] = (torch.randn(B * T, 3, nh, hd, generator=g0) * 1.5).bfloat16()
    qkv = qkv.view(B * T, 3 * nh * hs).to(DEV)
    scale = hd ** -0.5
    out = ops.attention_fwd(qkv, B, T, nh, hd, scale)
    assert out.shape == (B * T, nh * hd)
    x = qkv.view(B, T, 3, nh, hs)[..., :hd].float()
    q, k, v = [x[:, :, i].transpose(1, 2) for i in range(3)]
    ref = F.scaled_dot_product_attention(q, k, v, scale=scale).transpose(1, 2).reshape(B * T, nh * hd)
    close(out, ref, tol=1e-2)


def test_attention_peaked_rows_rescale():
    """Scores whose running max jumps by far more than the lazy-rescale threshold between kv tiles."""
    B, T, nh, hd = 1, 577, 2, 64
    g0 = torch.Generator().manual_seed(5)
    qkv = (torch.randn(B * T, 3, nh, hd, generator=g0)).bfloat16()
    qkv[:, 1] *= torch.linspace(0.1, 12.0, T).view(T, 1, 1).bfloat16()   # later keys score much higher
    qkv = qkv.view(B * T, 3 * nh * hd).to(DEV)
    out = ops.attention_fwd(qkv, B, T, nh, hd, 0.125)
    x = qkv.view(B, T, 3, nh, hd).float()
    q, k, v = [x[:, :, i].transpose(1, 2) for i in range(3)]
    ref = F.scaled_dot_product_attention(q, k, v, scale=0.125).transpose(1, 2).reshape(B * T, nh * hd)
    close(out, ref, tol=1e-2)


@pytest.mark.parametrize("T,hd,jump_at", [(577, 64, 256), (577, 64, 130), (729, 72, 480), (300, 64, 290)])
def test_attention_score_jump_beyond_the_stale_reference(T, hd, jump_at):
    """Online softmax under a sudden score jump.  The running reference max only moves when a kv tile's own max exceeds
    it by more than the lazy-rescale threshold (2^8); here the scores jump by > 100 log2 units at key `jump_at` (first
    image) and by a quarter of that (second image), in the middle of a kv tile and near the end of the row: the O
    accumulator and the row sum must be rescaled at that tile, and nothing may overflow.  (The name is historical: a
    'stale max' variant of the kernel, removed in round 3, motivated the case; the shipped kernel — head_dim 64 through
    attention3_sm100.cu, head_dim 72 through attention_sm100.cu — decides every tile's max from its own scores.)"""
    B, nh = 2, 2
    hs = ops.attn_head_stride(hd)
    g0 = torch.Generator().manual_seed(T + hd)
    x = torch.randn(B * T, 3, nh, hd, generator=g0) * 0.3
    x[:, 0] += 1.0                                            # q ~ ones
    t = torch.arange(B * T) % T
    big = (t >= jump_at).float().view(-1, 1, 1)
    x[:, 1] += big * 10.0                                     # k: +10 on every channel -> score jump ~ hd*10*scale
    x[B * T // 2:, 1] *= 0.25                                 # second image: a jump that stays below the redo threshold
    qkv = torch.zeros(B * T, 3, nh, hs, dtype=torch.bfloat16)
    qkv[..., :hd] = x.bfloat16()
    qkv = qkv.view(B * T, 3 * nh * hs).to(DEV)
    scale = hd ** -0.5
    out = ops.attention_fwd(qkv, B, T, nh, hd, scale)
    assert torch.isfinite(out.float()).all()
    xx = qkv.view(B, T, 3, nh, hs)[..., :hd].float()
    q, k, v = [xx[:, :, i].transpose(1, 2) for i in range(3)]
    ref = F.scaled_dot_product_attention(q, k, v, scale=scale).transpose(1, 2).reshape(B * T, nh * hd)
    close(out, ref, tol=1e-2)


def test_attention_rejects_unsupported_shapes():
    qkv = torch.zeros(8, 3 * 96, device=DEV).bfloat16()
    with pytest.raises(ops.MMError):
        ops.attention_fwd(qkv, 1, 8, 1, 96, 0.1)


def test_weighted_expert_sum():
    E, B, P, D = 5, 3, 7, 64
    x = torch.randn(E, B * P, D, device=DEV).bfloat16()
    w = torch.softmax(torch.randn(B, E, device=DEV), -1)
    perm = torch.tensor([2, 0, 4, 1, 3], dtype=torch.int32, device=DEV)
    for pm in (None, perm):
        out = ops.weighted_expert_sum(x, w, pm, B, P)
        ww = w if pm is None else w[:, pm.long()]
        st = x.view(E, B, P, D).permute(1, 0, 2, 3)  # [B,E,P,D]
        ref = (st * ww.to(torch.bfloat16)[:, :, None, None]).sum(1).reshape(B * P, D)
        close(out, ref, tol=1e-2)
