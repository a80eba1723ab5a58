"""torch.Tensor-level wrappers over the C-ABI (include/mm_b200.h).

PyTorch is plumbing here: it owns device memory and the CUDA stream; every op hands raw pointers
to the hand-written sm_100a kernels.  There is no fallback: a non-CUDA tensor raises.
"""
from __future__ import annotations

import torch

from . import _lib

MM_BF16, MM_F32 = 0, 1
ACT_NONE, ACT_QUICK_GELU, ACT_GELU_ERF, ACT_GELU_TANH, ACT_RELU = 0, 1, 2, 3, 4
ACT_BY_NAME = {
    None: ACT_NONE, "none": ACT_NONE, "quick_gelu": ACT_QUICK_GELU, "gelu": ACT_GELU_ERF,
    "gelu_pytorch_tanh": ACT_GELU_TANH, "gelu_tanh": ACT_GELU_TANH,
}
STATUS_INDEX_OOB, STATUS_COUNT_MISMATCH = 1, 2


class MMError(RuntimeError):
    pass


# -- instrumentation (bench.py): kernel-launch counter and optional per-op CUDA-event timing --------------------
KERNELS_PER_CALL = {"mm_attach_positions": 3, "mm_patch_embed": 2}
LAUNCHES = 0
_PROFILE = None  # dict: op name -> list[(start_event, end_event, work)] when enabled


def profile_start():
    global _PROFILE
    _PROFILE = {}


def profile_stop():
    """-> {op: (n_calls, total_ms, total_work)}; synchronises."""
    global _PROFILE
    prof, _PROFILE = _PROFILE, None
    torch.cuda.synchronize()
    return {k: (len(v), sum(a.elapsed_time(b) for a, b, _ in v), sum(w for _, _, w in v)) for k, v in (prof or {}).items()}


class _timed:
    def __init__(self, name, work=0.0):
        self.name, self.work = name, work

    def __enter__(self):
        global LAUNCHES
        LAUNCHES += KERNELS_PER_CALL.get(self.name, 1)
        if _PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *a):
        if _PROFILE is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _PROFILE.setdefault(self.name, []).append((self.e0, e1, self.work))
        return False


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return MM_BF16
    if t.dtype == torch.float32:
        return MM_F32
    raise TypeError(f"unsupported dtype {t.dtype} (bf16 / fp32 only)")


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise MMError("multimeditron_b200 ops need CUDA tensors (sm_100a kernels only, no CPU fallback)")


def _check(rc: int, what: str):
    if rc != 0:
        lib = _lib.load()
        names = {1: "MM_ERR_BAD_ARG", 2: "MM_ERR_CUDA", 3: "MM_ERR_UNSUPPORTED_ARCH"}
        raise MMError(f"{what} failed: {names.get(rc, rc)} (cuda error {lib.mm_last_cuda_error()})")


def new_status(device) -> torch.Tensor:
    return torch.zeros(1, dtype=torch.int32, device=device)


def raise_for_status(status: torch.Tensor):
    """Lazy check of the device status word (one D2H sync). Raises what the reference raises."""
    s = int(status.item())
    if s & STATUS_INDEX_OOB:
        raise IndexError("index out of range in splice (batch_idx/token_range/input_ids)")
    if s & STATUS_COUNT_MISMATCH:
        # model.py:441 -> index_put_ shape mismatch is a RuntimeError in the reference
        raise RuntimeError("shape mismatch: number of attachment tokens != number of modality embeddings")


# ------------------------------------------------------------------------------------------ splice
def attach_positions(input_ids: torch.Tensor, attachment_id: int, expected_rows: int = -1,
                     want_src_of_pos: bool = True, status: torch.Tensor | None = None):
    """-> (dest_rows int32[cap], src_of_pos int32[n_ids] | None, count int32[1]) without host sync."""
    _cuda(input_ids)
    lib = _lib.load()
    ids = input_ids.contiguous().view(-1)
    assert ids.dtype == torch.int64
    n = ids.numel()
    cap = expected_rows if expected_rows >= 0 else n
    dev = ids.device
    dest_rows = torch.empty(max(cap, 1), dtype=torch.int32, device=dev)
    sop = torch.empty(max(n, 1), dtype=torch.int32, device=dev) if want_src_of_pos else None
    count = torch.empty(1, dtype=torch.int32, device=dev)
    ws = torch.empty(lib.mm_attach_positions_workspace_bytes(n), dtype=torch.uint8, device=dev)
    with _timed("mm_attach_positions", float(n) * 8):
        rc = lib.mm_attach_positions(_ptr(ids), n, int(attachment_id), _ptr(dest_rows), cap, _ptr(sop), _ptr(count),
                                 int(expected_rows), _ptr(status), _ptr(ws), _stream())
    _check(rc, "mm_attach_positions")
    return dest_rows[:cap], (sop[:n] if sop is not None else None), count


def splice_scatter(dest: torch.Tensor, src: torch.Tensor, dest_rows: torch.Tensor,
                   status: torch.Tensor | None = None) -> torch.Tensor:
    """dest.view(-1,H)[dest_rows[k]] = src.view(-1,H)[k]   (in place; model.py:441-444)."""
    _cuda(dest, src, dest_rows)
    lib = _lib.load()
    H = dest.shape[-1]
    assert dest.is_contiguous() and src.is_contiguous() and src.shape[-1] == H
    assert dest_rows.dtype == torch.int32 and dest_rows.is_contiguous()
    n_rows = src.numel() // H if H else 0
    assert dest_rows.numel() >= n_rows
    with _timed("mm_splice_scatter", 2.0 * n_rows * H * src.element_size()):
        rc = lib.mm_splice_scatter(_ptr(dest), dest.numel() // H if H else 0, H, _dt(dest), _ptr(src), n_rows, _dt(src),
                               _ptr(dest_rows), _ptr(status), _stream())
    _check(rc, "mm_splice_scatter")
    return dest


def splice_scatter_idx(dest: torch.Tensor, src: torch.Tensor, batch_idx: torch.Tensor, token_range: torch.Tensor,
                       status: torch.Tensor | None = None) -> torch.Tensor:
    """dest[batch_idx, token_range] = src.view(-1,H)  with the collator's int64 index vectors."""
    _cuda(dest, src, batch_idx, token_range)
    lib = _lib.load()
    assert dest.dim() == 3 and dest.is_contiguous() and src.is_contiguous()
    B, S, H = dest.shape
    n_rows = src.numel() // H if H else 0
    if batch_idx.numel() != token_range.numel() or batch_idx.numel() != n_rows:
        # same class/condition as index_put_ in the reference (host-known sizes -> no device round trip)
        raise RuntimeError(
            f"shape mismatch: value tensor of shape [{n_rows}, {H}] cannot be broadcast to indexing result of "
            f"shape [{batch_idx.numel()}, {H}]")
    bi = batch_idx.contiguous().to(torch.int64)
    tr = token_range.contiguous().to(torch.int64)
    with _timed("mm_splice_scatter_idx", 2.0 * n_rows * H * src.element_size()):
        rc = lib.mm_splice_scatter_idx(_ptr(dest), B, S, H, _dt(dest), _ptr(src), n_rows, _dt(src), _ptr(bi), _ptr(tr),
                                   _ptr(status), _stream())
    _check(rc, "mm_splice_scatter_idx")
    return dest


def embed_gather(table: torch.Tensor, input_ids: torch.Tensor, status: torch.Tensor | None = None,
                 skip_id: int | None = None) -> torch.Tensor:
    """table[input_ids] (model.py:433).  skip_id: rows whose id equals it are NOT written — for callers that splice
    every attachment position afterwards (the rows of `out` at those positions are uninitialised until then)."""
    _cuda(table, input_ids)
    lib = _lib.load()
    V, H = table.shape
    ids = input_ids.contiguous()
    out = torch.empty(*ids.shape, H, dtype=table.dtype, device=table.device)
    n_copy = ids.numel()
    with _timed("mm_embed_gather", 2.0 * n_copy * H * table.element_size()):
        if skip_id is None:
            rc = lib.mm_embed_gather(_ptr(out), ids.numel(), H, _dt(table), _ptr(table.contiguous()), V, _ptr(ids),
                                     _ptr(status), _stream())
        else:
            rc = lib.mm_embed_gather_skip(_ptr(out), ids.numel(), H, _dt(table), _ptr(table.contiguous()), V, _ptr(ids),
                                          int(skip_id), _ptr(status), _stream())
    _check(rc, "mm_embed_gather")
    return out


def gather_rows(src: torch.Tensor, rows: torch.Tensor, out_dtype=None, status: torch.Tensor | None = None) -> torch.Tensor:
    """out[k] = src.view(-1,H)[rows[k]]  (rows int32) — the splice's backward (gradient of the projected tokens)."""
    _cuda(src, rows)
    lib = _lib.load()
    H = src.shape[-1]
    s2 = src.contiguous().view(-1, H)
    assert rows.dtype == torch.int32 and rows.is_contiguous()
    out = torch.empty(rows.numel(), H, dtype=out_dtype or src.dtype, device=src.device)
    with _timed("mm_gather_rows", 2.0 * rows.numel() * H * src.element_size()):
        rc = lib.mm_gather_rows(_ptr(out), rows.numel(), H, _dt(out), _ptr(s2), s2.shape[0], _dt(s2), _ptr(rows),
                                _ptr(status), _stream())
    _check(rc, "mm_gather_rows")
    return out


def embed_splice_fused(table: torch.Tensor, input_ids: torch.Tensor, src_of_pos: torch.Tensor, src: torch.Tensor,
                       status: torch.Tensor | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
    _cuda(table, input_ids, src_of_pos, src)
    lib = _lib.load()
    V, H = table.shape
    ids = input_ids.contiguous()
    if out is None:
        out = torch.empty(*ids.shape, H, dtype=table.dtype, device=table.device)
    n_rows = src.numel() // H if H else 0
    with _timed("mm_embed_splice_fused", 2.0 * ids.numel() * H * out.element_size()):
        rc = lib.mm_embed_splice_fused(_ptr(out), ids.numel(), H, _dt(out), _ptr(table.contiguous()), V, _ptr(ids),
                                   _ptr(src_of_pos), _ptr(src.contiguous()), n_rows, _dt(src), _ptr(status),
                                   _stream())
    _check(rc, "mm_embed_splice_fused")
    return out


# ------------------------------------------------------------------------------------------ ViT blocks
def layernorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float, out: torch.Tensor | None = None,
              rows_in_group: int = 0, rows_out_group: int = 0, in_row_offset: int = 0) -> torch.Tensor:
    _cuda(x, gamma, beta)
    lib = _lib.load()
    D = x.shape[-1]
    assert x.dtype == torch.bfloat16 and x.is_contiguous()
    if out is None:
        out = torch.empty_like(x)
    M = out.numel() // D
    with _timed("mm_layernorm", 4.0 * M * D):
        rc = lib.mm_layernorm(_ptr(x), _ptr(out), _ptr(gamma), _ptr(beta), M, D, float(eps), rows_in_group,
                          rows_out_group, in_row_offset, _stream())
    _check(rc, "mm_layernorm")
    return out


def gemm_bias_act(a: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, act: int = ACT_NONE,
                  residual: torch.Tensor | None = None, res_mod: int = 0, res_row_off: int = 0,
                  out: torch.Tensor | None = None, out_row_map: torch.Tensor | None = None) -> torch.Tensor:
    """out = act(a @ w.T + bias) [+ residual]; a [M,K], w [N,K] bf16 (row strides may exceed K)."""
    _cuda(a, w, bias, residual, out, out_row_map)
    lib = _lib.load()
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and a.dim() == 2 and w.dim() == 2
    assert a.stride(1) == 1 and w.stride(1) == 1
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K
    if out is None:
        assert out_row_map is None
        out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
    assert out.dtype == torch.bfloat16 and out.stride(-1) == 1
    ldc = out.stride(-2) if out.dim() >= 2 else N
    ldr = residual.stride(-2) if residual is not None else 0
    n_out_rows = (out.numel() // N if out.is_contiguous() else out.shape[0]) if out_row_map is not None else 0
    with _timed("mm_gemm_bias_act", 2.0 * M * N * K):
        rc = lib.mm_gemm_bias_act(_ptr(a), a.stride(0), _ptr(w), w.stride(0), _ptr(bias), _ptr(out), ldc, M, N, K,
                              int(act), _ptr(residual), ldr, int(res_mod), int(res_row_off), _ptr(out_row_map),
                                  n_out_rows, _stream())
    _check(rc, "mm_gemm_bias_act")
    return out


def gemm_ln_bias_act(a: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, act: int = ACT_NONE,
                     residual: torch.Tensor | None = None, out: torch.Tensor | None = None,
                     out_row_map: torch.Tensor | None = None, stats_out: torch.Tensor | None = None,
                     ln_stats: torch.Tensor | None = None, eps: float = 1e-5) -> torch.Tensor:
    """mm_gemm_ln_bias_act (include/mm_b200.h): GEMM that emits row statistics of what it stores (stats_out
    fp32 [rows, slots, 2], zero-initialised once) and/or consumes them (ln_stats) to apply a folded LayerNorm to
    its A operand; in that case (w, bias) must come from `fold_layernorm`."""
    _cuda(a, w, bias, residual, out, out_row_map, stats_out, ln_stats)
    lib = _lib.load()
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and a.dim() == 2 and w.dim() == 2
    assert a.stride(1) == 1 and w.stride(1) == 1
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K
    if out is None:
        assert out_row_map is None
        out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
    assert out.dtype == torch.bfloat16 and out.stride(-1) == 1
    ldc = out.stride(-2) if out.dim() >= 2 else N
    ldr = residual.stride(-2) if residual is not None else 0
    n_out_rows = (out.numel() // N if out.is_contiguous() else out.shape[0]) if out_row_map is not None else 0
    if stats_out is not None:
        rows = n_out_rows if out_row_map is not None else M
        assert stats_out.dtype == torch.float32 and stats_out.is_contiguous()
        assert stats_out.numel() >= rows * row_stats_slots(N) * 2
    if ln_stats is not None:
        assert ln_stats.dtype == torch.float32 and ln_stats.is_contiguous()
        assert ln_stats.numel() >= M * row_stats_slots(K) * 2
    with _timed("mm_gemm_bias_act", 2.0 * M * N * K):
        rc = lib.mm_gemm_ln_bias_act(_ptr(a), a.stride(0), _ptr(w), w.stride(0), _ptr(bias), _ptr(out), ldc, M, N, K,
                                     int(act), _ptr(residual), ldr, _ptr(out_row_map), n_out_rows, _ptr(stats_out),
                                     _ptr(ln_stats), float(eps), _stream())
    _check(rc, "mm_gemm_ln_bias_act")
    return out


def gemm_grouped(a: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, groups: int, act: int = ACT_NONE,
                 residual: torch.Tensor | None = None, out: torch.Tensor | None = None,
                 out_row_map: torch.Tensor | None = None, stats_out: torch.Tensor | None = None,
                 ln_stats: torch.Tensor | None = None, eps: float = 1e-5) -> torch.Tensor:
    """`groups` same-shaped GEMMs in ONE launch (mm_gemm_grouped): a [groups*M, K], w [groups*N, K], bias [groups*N]
    -> out [groups*M, N]; group g's rows multiply w[g*N:(g+1)*N].  M must be a multiple of 256."""
    _cuda(a, w, bias, residual, out, out_row_map, stats_out, ln_stats)
    lib = _lib.load()
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and a.dim() == 2 and w.dim() == 2
    assert a.stride(1) == 1 and w.stride(1) == 1
    Mt, K = a.shape
    assert Mt % groups == 0 and w.shape[0] % groups == 0 and w.shape[1] == K
    M, N = Mt // groups, w.shape[0] // groups
    if out is None:
        assert out_row_map is None
        out = torch.empty(Mt, N, dtype=torch.bfloat16, device=a.device)
    assert out.dtype == torch.bfloat16 and out.stride(-1) == 1
    ldc = out.stride(-2) if out.dim() >= 2 else N
    ldr = residual.stride(-2) if residual is not None else 0
    n_out_rows = (out.numel() // N if out.is_contiguous() else out.shape[0]) if out_row_map is not None else 0
    with _timed("mm_gemm_bias_act", 2.0 * Mt * N * K):
        rc = lib.mm_gemm_grouped(_ptr(a), a.stride(0), _ptr(w), w.stride(0), _ptr(bias), _ptr(out), ldc, M, N, K,
                                 int(groups), int(act), _ptr(residual), ldr, _ptr(out_row_map), n_out_rows,
                                 _ptr(stats_out), _ptr(ln_stats), float(eps), _stream())
    _check(rc, "mm_gemm_grouped")
    return out


GEMM_FLAG_1CTA = 0x100   # per-call: OR into `act` to run the 1-CTA kernel instead of the CTA pair (bit-identical results)


def row_stats_slots(n: int) -> int:
    return int(_lib.load().mm_gemm_row_stats_slots(n))


def fold_layernorm(w: torch.Tensor, bias: torch.Tensor | None, gamma: torch.Tensor, beta: torch.Tensor):
    """(W'', b') for the consumer side of mm_gemm_ln_bias_act — one launch of mm_fold_layernorm (load time).
    W'' = gamma*W with each row re-centred to zero sum, rounded to bf16 with error feedback along k (sigma-delta), so
    the bf16 row sums stay within half an ulp of zero and the LayerNorm mean term cancels inside the MMA;
    b' = bias + W @ beta, rounded to bf16."""
    _cuda(w, bias, gamma, beta)
    lib = _lib.load()
    assert w.dtype == torch.bfloat16 and w.dim() == 2 and w.stride(1) == 1
    N, K = w.shape
    g, b = gamma.to(torch.bfloat16).contiguous(), beta.to(torch.bfloat16).contiguous()
    bi = bias.to(torch.bfloat16).contiguous() if bias is not None else None
    q = torch.empty(N, K, dtype=torch.bfloat16, device=w.device)
    bo = torch.empty(N, dtype=torch.bfloat16, device=w.device)
    with _timed("mm_fold_layernorm", 4.0 * N * K):
        rc = lib.mm_fold_layernorm(_ptr(w), w.stride(0), _ptr(bi), _ptr(g), _ptr(b), _ptr(q), K, _ptr(bo), N, K, _stream())
    _check(rc, "mm_fold_layernorm")
    return q, bo


_patch_ws: dict = {}   # (device, stream) -> scratch for the staged patch rows (grown on demand, reused across calls)


def patch_embed(pixels: torch.Tensor, weight_padded: torch.Tensor, conv_bias, pos_emb: torch.Tensor, cls,
                patch: int, out: torch.Tensor | None = None, staged: bool = True) -> torch.Tensor:
    """pixels [B,3,img,img] fp32/bf16; weight_padded [D, Kpad] bf16 (zero padded) -> [B,T,D] bf16.
    staged=True: patch rows -> bf16 scratch -> CTA-pair GEMM (mm_patch_embed_ws); False: in-kernel gather."""
    _cuda(pixels, weight_padded, conv_bias, pos_emb, cls)
    lib = _lib.load()
    B, Cc, img, img2 = pixels.shape
    assert Cc == 3 and img == img2 and pixels.is_contiguous()
    D = weight_padded.shape[0]
    T = pos_emb.shape[0]
    if out is None:
        out = torch.empty(B, T, D, dtype=torch.bfloat16, device=pixels.device)
    with _timed("mm_patch_embed", 2.0 * B * (T - (1 if cls is not None else 0)) * D * 3 * patch * patch):
        if staged and B > 0:
            global LAUNCHES
            LAUNCHES += 1   # patch_rows_kernel (the context manager counts the GEMM and the CLS kernel)
            need = int(lib.mm_patch_embed_workspace_bytes(B, img, patch))
            key = (pixels.device, _stream())     # one scratch per (device, stream): concurrent streams never share it
            ws = _patch_ws.get(key)
            if ws is None or ws.numel() < need:
                ws = _patch_ws[key] = torch.empty(need, dtype=torch.uint8, device=pixels.device)
            rc = lib.mm_patch_embed_ws(_ptr(pixels), _dt(pixels), B, img, patch, _ptr(weight_padded),
                                       weight_padded.stride(0), _ptr(conv_bias), _ptr(pos_emb), _ptr(cls), _ptr(out), D,
                                       _ptr(ws), ws.numel(), _stream())
        else:
            rc = lib.mm_patch_embed(_ptr(pixels), _dt(pixels), B, img, patch, _ptr(weight_padded),
                                    weight_padded.stride(0), _ptr(conv_bias), _ptr(pos_emb), _ptr(cls), _ptr(out), D,
                                    _stream())
    _check(rc, "mm_patch_embed")
    return out


def attn_head_stride(head_dim: int) -> int:
    """Column slot of one head in the qkv buffer (64 -> 64, 72 -> 80 with 8 zero pad columns)."""
    hs = _lib.load().mm_attention_head_stride(head_dim)
    if hs < 0:
        raise MMError(f"mm_attention_fwd: head_dim {head_dim} is not supported (64 or 72)")
    return int(hs)


def attention_fwd(qkv: torch.Tensor, B: int, T: int, heads: int, head_dim: int, scale: float,
                  out: torch.Tensor | None = None) -> torch.Tensor:
    """qkv [B*T, 3*heads*attn_head_stride(head_dim)] bf16 -> out [B*T, heads*head_dim] bf16."""
    _cuda(qkv)
    lib = _lib.load()
    hs = attn_head_stride(head_dim)
    D = heads * head_dim
    assert qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and qkv.numel() == B * T * 3 * heads * hs
    if out is None:
        out = torch.empty(B * T, D, dtype=torch.bfloat16, device=qkv.device)
    assert out.dtype == torch.bfloat16 and out.is_contiguous() and out.numel() == B * T * D
    with _timed("mm_attention_fwd", 4.0 * B * heads * T * T * head_dim):
        rc = lib.mm_attention_fwd(_ptr(qkv), _ptr(out), B, T, heads, head_dim, float(scale), _stream())
    _check(rc, "mm_attention_fwd")
    return out


def weighted_expert_sum(x: torch.Tensor, w: torch.Tensor, perm: torch.Tensor | None, B: int, P: int) -> torch.Tensor:
    """x [E, B*P, D] bf16, w [B,E] fp32, perm int32[E] | None -> [B*P, D] bf16."""
    _cuda(x, w, perm)
    lib = _lib.load()
    E, R, D = x.shape
    assert R == B * P and x.is_contiguous() and w.dtype == torch.float32 and w.is_contiguous()
    out = torch.empty(R, D, dtype=torch.bfloat16, device=x.device)
    with _timed("mm_weighted_expert_sum", 2.0 * (E + 1) * R * D):
        rc = lib.mm_weighted_expert_sum(_ptr(x), _ptr(w), _ptr(perm), _ptr(out), E, B, P, D, _stream())
    _check(rc, "mm_weighted_expert_sum")
    return out


# --------------------------------------------------------------------------------- ResNet gate: row staging / pooling
def conv_out_size(n: int, k: int, stride: int, pad: int) -> int:
    return (n + 2 * pad - k) // stride + 1


def conv_rows_nhwc(x: torch.Tensor, kh: int, kw: int, stride: int, pad: int, kpad: int) -> torch.Tensor:
    """x [B,H,W,C] bf16 -> [B*Ho*Wo, kpad] bf16 rows of a kh x kw convolution (k = (ky*kw+kx)*C + c)."""
    _cuda(x)
    B, H, W, Cc = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous()
    Ho, Wo = conv_out_size(H, kh, stride, pad), conv_out_size(W, kw, stride, pad)
    out = torch.empty(B * Ho * Wo, kpad, dtype=torch.bfloat16, device=x.device)
    with _timed("mm_conv_rows", 2.0 * out.numel() * 2):
        rc = _lib.load().mm_conv_rows_nhwc(_ptr(x), _ptr(out), B, H, W, Cc, kh, kw, stride, pad, kpad, _stream())
    _check(rc, "mm_conv_rows_nhwc")
    return out


def conv_rows_stem(x: torch.Tensor, kh: int, kw: int, stride: int, pad: int, kpad: int) -> torch.Tensor:
    """x [B,3,H,W] fp32|bf16 (pixel_values) -> [B*Ho*Wo, kpad] bf16 rows of the stem convolution."""
    _cuda(x)
    B, Cc, H, W = x.shape
    assert Cc == 3 and x.is_contiguous()
    Ho, Wo = conv_out_size(H, kh, stride, pad), conv_out_size(W, kw, stride, pad)
    out = torch.empty(B * Ho * Wo, kpad, dtype=torch.bfloat16, device=x.device)
    with _timed("mm_conv_rows", float(x.numel() * x.element_size() + out.numel() * 2)):
        rc = _lib.load().mm_conv_rows_stem(_ptr(x), _dt(x), _ptr(out), B, H, W, kh, kw, stride, pad, kpad, _stream())
    _check(rc, "mm_conv_rows_stem")
    return out


def maxpool_nhwc(x: torch.Tensor) -> torch.Tensor:
    _cuda(x)
    B, H, W, Cc = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous()
    out = torch.empty(B, conv_out_size(H, 3, 2, 1), conv_out_size(W, 3, 2, 1), Cc, dtype=torch.bfloat16, device=x.device)
    with _timed("mm_pool", float(x.numel() * 2 + out.numel() * 2)):
        rc = _lib.load().mm_maxpool_nhwc(_ptr(x), _ptr(out), B, H, W, Cc, _stream())
    _check(rc, "mm_maxpool_nhwc")
    return out


def avgpool_nhwc(x: torch.Tensor) -> torch.Tensor:
    """[B,HW,C] bf16 -> [B,C] bf16."""
    _cuda(x)
    B, HW, Cc = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous()
    out = torch.empty(B, Cc, dtype=torch.bfloat16, device=x.device)
    with _timed("mm_pool", float(x.numel() * 2)):
        rc = _lib.load().mm_avgpool_nhwc(_ptr(x), _ptr(out), B, HW, Cc, _stream())
    _check(rc, "mm_avgpool_nhwc")
    return out


def cross_attention(q: torch.Tensor, kv: torch.Tensor, heads: int, scale: float | None = None) -> torch.Tensor:
    """q [B,Nq,C] bf16, kv [B,Nkv,2C] bf16 ([K | V] columns) -> [B,Nq,C] bf16 (attention.py:48-101, eval, no mask)."""
    _cuda(q, kv)
    lib = _lib.load()
    B, Nq, Cc = q.shape
    assert kv.shape[0] == B and kv.shape[2] == 2 * Cc and q.is_contiguous() and kv.is_contiguous()
    assert q.dtype == torch.bfloat16 and kv.dtype == torch.bfloat16
    Nkv = kv.shape[1]
    out = torch.empty_like(q)
    sc = float(scale) if scale is not None else (Cc // heads) ** -0.5
    with _timed("mm_cross_attention", 4.0 * B * Nq * Nkv * Cc):
        rc = lib.mm_cross_attention(_ptr(q), _ptr(kv), _ptr(out), B, Nq, Nkv, int(heads), Cc, sc, _stream())
    _check(rc, "mm_cross_attention")
    return out


def position_ids(attention_mask: torch.Tensor, input_ids: torch.Tensor | None = None, attachment_id: int = -1,
                 tokens_per_image: int = 0, grid: int = 0, use_2d: bool = False,
                 status: torch.Tensor | None = None) -> torch.Tensor:
    """data_loader.py:155-188 on the device: [B,S] int64 (1-D) or [B,S,2] int64 (2-D image position ids)."""
    _cuda(attention_mask, input_ids, status)
    lib = _lib.load()
    B, S = attention_mask.shape
    m = attention_mask.contiguous()
    if m.dtype == torch.bool:
        m = m.view(torch.uint8)
    if m.dtype not in (torch.int64, torch.uint8):
        raise TypeError(f"attention_mask must be int64, uint8 or bool, got {m.dtype}")
    ids = None
    if use_2d:
        assert input_ids is not None and input_ids.dtype == torch.int64 and input_ids.shape == attention_mask.shape
        ids = input_ids.contiguous()
    out = torch.empty((B, S, 2) if use_2d else (B, S), dtype=torch.int64, device=m.device)
    with _timed("mm_position_ids", float(B) * S * (8 + 8 * (2 if use_2d else 1))):
        rc = lib.mm_position_ids(_ptr(ids), _ptr(m), int(m.dtype == torch.uint8), B, S, int(attachment_id),
                                 int(tokens_per_image), int(grid), int(use_2d), _ptr(out), _ptr(status), _stream())
    _check(rc, "mm_position_ids")
    return out


# --------------------------------------------------------------------------------- projector training step
def gelu_fwd(z: torch.Tensor) -> torch.Tensor:
    _cuda(z)
    assert z.dtype == torch.bfloat16 and z.is_contiguous()
    a = torch.empty_like(z)
    with _timed("mm_gelu_fwd", 4.0 * z.numel()):
        rc = _lib.load().mm_gelu_fwd(_ptr(z), _ptr(a), z.numel(), _stream())
    _check(rc, "mm_gelu_fwd")
    return a


def gelu_bwd_colsum(da: torch.Tensor, z: torch.Tensor | None):
    """-> (dz bf16 [M,N] | None, bias_grad fp32 [N]);  z None: column sum of `da` only."""
    _cuda(da, z)
    lib = _lib.load()
    assert da.dtype == torch.bfloat16 and da.is_contiguous() and da.dim() == 2
    M, N = da.shape
    dz = None
    if z is not None:
        assert z.shape == da.shape and z.dtype == torch.bfloat16 and z.is_contiguous()
        dz = torch.empty_like(da)
    bg = torch.empty(N, dtype=torch.float32, device=da.device)
    ws = torch.empty(max(256, int(lib.mm_colsum_workspace_bytes(M, N))), dtype=torch.uint8, device=da.device)
    global LAUNCHES
    LAUNCHES += 1
    with _timed("mm_gelu_bwd_colsum", (6.0 if z is not None else 2.0) * M * N):
        rc = lib.mm_gelu_bwd_colsum(_ptr(da), _ptr(z), _ptr(dz), _ptr(bg), _ptr(ws), M, N, _stream())
    _check(rc, "mm_gelu_bwd_colsum")
    return dz, bg


def transpose(x: torch.Tensor) -> torch.Tensor:
    """[R, C] bf16 -> contiguous [C, R]."""
    _cuda(x)
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and x.dim() == 2
    R, C = x.shape
    out = torch.empty(C, R, dtype=torch.bfloat16, device=x.device)
    with _timed("mm_transpose", 4.0 * R * C):
        rc = _lib.load().mm_transpose(_ptr(x), _ptr(out), R, C, _stream())
    _check(rc, "mm_transpose")
    return out


# --------------------------------------------------------------------------------- image pre-processing
def preprocess_images(src: torch.Tensor, desc: torch.Tensor, size: int, kmax: int, max_h: int, max_w: int,
                      tmp_bytes: int, mean255, std255, out_dtype=torch.float32) -> torch.Tensor:
    """src uint8 device bytes (packed HWC images), desc int64 [B,8] device (include/mm_b200.h) -> [B,3,size,size]."""
    import ctypes as C
    _cuda(src, desc)
    lib = _lib.load()
    B = desc.shape[0]
    assert src.dtype == torch.uint8 and desc.dtype == torch.int64 and desc.is_contiguous() and desc.shape[1] == 8
    out = torch.empty(B, 3, size, size, dtype=out_dtype, device=src.device)
    ws = torch.empty(max(256, lib.mm_preprocess_workspace_bytes(B, size, kmax)), dtype=torch.uint8, device=src.device)
    tmp = torch.empty(max(256, tmp_bytes), dtype=torch.uint8, device=src.device)
    m = (C.c_float * 3)(*[float(v) for v in mean255])
    s = (C.c_float * 3)(*[float(v) for v in std255])
    global LAUNCHES
    LAUNCHES += 2  # three kernels per call (the context manager counts one)
    with _timed("mm_preprocess_images", float(src.numel()) + out.numel() * out.element_size()):
        rc = lib.mm_preprocess_images(_ptr(src), _ptr(desc), B, size, kmax, max_h, max_w, _ptr(ws), _ptr(tmp), m, s,
                                      _ptr(out), _dt(out), _stream())
    _check(rc, "mm_preprocess_images")
    return out
