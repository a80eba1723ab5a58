"""Training step of the projector on the sm_100a kernels.

The reference's ALIGNMENT stage (`MultiModalModelForCausalLM.freeze_for_alignment`, model/model.py:310-326; caller
train/trainer.py:135-144) trains ONLY `MLPProjector` (model/projectors/mlp.py:33-39:
Linear(D,D) -> GELU -> Linear(D,H) -> GELU -> Linear(H,H)) with the image tower and the LLM frozen.  `ProjectorFunction`
is that module's forward + backward as one `torch.autograd.Function`:

  forward   z0 = x W0ᵀ + b0, a0 = gelu(z0), z1 = a0 W2ᵀ + b2, a1 = gelu(z1), y = a1 W4ᵀ + b4       (bf16, fp32 accumulate;
            every intermediate rounded to bf16 exactly where nn.Linear / nn.GELU round in the pure-bf16 regime)
  backward  dWi = dZiᵀ · A(i-1)     (wgrad: tcgen05 GEMM, K = number of tokens, operands transposed by mm_transpose)
            dA(i-1) = dZi · Wi      (dgrad: tcgen05 GEMM on Wiᵀ)
            dZi = dAi ⊙ gelu'(Zi), dbi = Σ_rows dZi    (mm_gelu_bwd_colsum, deterministic)

`SpliceRowsFunction` makes the splice statement (model.py:441-444) differentiable: forward = in-place row scatter,
backward = row gather of the incoming gradient at the same positions.
There is no PyTorch fallback: every op is a C-ABI call; autograd only sequences them.
"""
from __future__ import annotations

import torch

from . import ops


def _bf(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.bfloat16).contiguous()


class ProjectorFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w0, b0, w2, b2, w4, b4):
        xb = _bf(x)
        M = xb.shape[0]
        Mp = (M + 7) // 8 * 8      # the wgrad GEMMs contract over the token dimension: K must be a multiple of 8.
        if Mp != M:                # Zero rows are appended (their dy is zero in backward: no contribution anywhere)
            xp = torch.zeros(Mp, xb.shape[1], dtype=xb.dtype, device=xb.device)
            xp[:M] = xb
            xb = xp
        ctx.rows = M
        W0, B0, W2, B2, W4, B4 = map(_bf, (w0, b0, w2, b2, w4, b4))
        z0 = ops.gemm_bias_act(xb, W0, B0)
        a0 = ops.gelu_fwd(z0)
        z1 = ops.gemm_bias_act(a0, W2, B2)
        a1 = ops.gelu_fwd(z1)
        y = ops.gemm_bias_act(a1, W4, B4)
        ctx.save_for_backward(xb, z0, a0, z1, a1, W0, W2, W4)
        ctx.dtypes = (x.dtype, w0.dtype, b0.dtype, w2.dtype, b2.dtype, w4.dtype, b4.dtype)
        return y[:M] if Mp != M else y

    @staticmethod
    def backward(ctx, dy):
        xb, z0, a0, z1, a1, W0, W2, W4 = ctx.saved_tensors
        need = ctx.needs_input_grad
        dyb = _bf(dy)
        M = ctx.rows
        if xb.shape[0] != M:
            dp = torch.zeros(xb.shape[0], dyb.shape[1], dtype=dyb.dtype, device=dyb.device)
            dp[:M] = dyb
            dyb = dp

        def wgrad(dz, a):        # [N, K] = dzᵀ[N, M] · a[M, K]  ->  GEMM(A = dzᵀ, "W" = aᵀ)
            return ops.gemm_bias_act(ops.transpose(dz), ops.transpose(a))

        def dgrad(dz, W):        # [M, K] = dz[M, N] · W[N, K]   ->  GEMM(A = dz, "W" = Wᵀ)
            return ops.gemm_bias_act(dz, ops.transpose(W))

        _, db4 = ops.gelu_bwd_colsum(dyb, None)
        dW4 = wgrad(dyb, a1) if need[5] else None
        da1 = dgrad(dyb, W4)
        dz1, db2 = ops.gelu_bwd_colsum(da1, z1)
        dW2 = wgrad(dz1, a0) if need[3] else None
        da0 = dgrad(dz1, W2)
        dz0, db0 = ops.gelu_bwd_colsum(da0, z0)
        dW0 = wgrad(dz0, xb) if need[1] else None
        dx = dgrad(dz0, W0) if need[0] else None
        if dx is not None and dx.shape[0] != M:
            dx = dx[:M]
        dts = ctx.dtypes
        cast = lambda g, dt: None if g is None else g.to(dt)  # noqa: E731
        return (cast(dx, dts[0]), cast(dW0, dts[1]), cast(db0, dts[2]) if need[2] else None, cast(dW2, dts[3]),
                cast(db2, dts[4]) if need[4] else None, cast(dW4, dts[5]), cast(db4, dts[6]) if need[6] else None)


def projector_forward(feats: torch.Tensor, projector: torch.nn.Module) -> torch.Tensor:
    """`projector` is a reference MLPProjector (or anything with .projection[0|2|4] Linear layers)."""
    p = projector.projection
    return ProjectorFunction.apply(feats, p[0].weight, p[0].bias, p[2].weight, p[2].bias, p[4].weight, p[4].bias)


class SpliceRowsFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dest2d, src, rows, status):
        ops.splice_scatter(dest2d, src.detach().contiguous(), rows, status=status)
        ctx.mark_dirty(dest2d)
        ctx.save_for_backward(rows)
        ctx.src_dtype = src.dtype
        return dest2d

    @staticmethod
    def backward(ctx, grad):
        (rows,) = ctx.saved_tensors
        g = grad.contiguous()
        gsrc = ops.gather_rows(g, rows, out_dtype=g.dtype).to(ctx.src_dtype) if ctx.needs_input_grad[1] else None
        # index_put_ semantics for the base tensor: the rows this splice overwrote carry no gradient back to what was
        # there before, every other row passes through.  The base needs that gradient when it is itself the output of
        # an earlier splice whose source trains (two modality types with trainable projectors); a base gathered from
        # the frozen table does not require grad and nothing is materialised for it.
        gdest = None
        if ctx.needs_input_grad[0]:
            gdest = g.clone()
            zeros = torch.zeros(rows.numel(), g.shape[-1], dtype=g.dtype, device=g.device)
            ops.splice_scatter(gdest, zeros, rows)       # status=None: rows < 0 ("skip") are ignored, as in forward
        return gdest, gsrc, None, None


def splice_rows(dest2d: torch.Tensor, src: torch.Tensor, rows: torch.Tensor, status=None) -> torch.Tensor:
    """dest2d[rows[k]] = src[k] in place, differentiable w.r.t. src (model.py:441-444)."""
    return SpliceRowsFunction.apply(dest2d, src, rows.to(torch.int32).contiguous(), status)


class WeightedExpertSumFunction(torch.autograd.Function):
    """fused[b,p,:] = sum_e w[b, perm[e]] * x[e,b,p,:]  (image_modality_moe.py:171-177, image_modality_moe_pep.py:212-216)
    with a backward for x (the per-expert projector outputs of the PEP variant, whose projectors train in the ALIGNMENT
    stage): dx[e] = w[:, perm[e]] * dfused.  The gate is frozen there (freeze_modality_embedder), so w gets no gradient."""

    @staticmethod
    def forward(ctx, x, w, perm, B, P):
        ctx.save_for_backward(w, perm if perm is not None else torch.empty(0))
        ctx.meta = (B, P, perm is not None, x.shape[0])
        return ops.weighted_expert_sum(x.detach().contiguous(), w, perm, B, P)

    @staticmethod
    def backward(ctx, g):
        w, perm = ctx.saved_tensors
        B, P, has_perm, E = ctx.meta
        gb = g.detach().to(torch.bfloat16).contiguous()[None]          # [1, B*P, D]
        outs = []
        for e in range(E):
            col = int(perm[e]) if has_perm else e
            outs.append(ops.weighted_expert_sum(gb, w[:, col:col + 1].contiguous(), None, B, P))   # bf16(w * g), like autograd
        return torch.stack(outs, 0), None, None, None, None


def weighted_expert_sum(x, w, perm, B, P):
    return WeightedExpertSumFunction.apply(x, w, perm, B, P)
