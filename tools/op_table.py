"""Per-op table: every kernel of the path next to the library op PyTorch 2.11 dispatches for the same work on the same
B200 ("the bar is cuBLASLt/cuDNN/flash-attn as PyTorch dispatches them").  CUDA-event timed, 3 warm-ups,
inputs larger than L2 or rotated; prints one JSON object per line and a summary table.
    python tools/op_table.py [--images 256] > profiles/rNN_op_table.txt
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multimeditron_b200 import ops  # noqa: E402
from multimeditron_b200.configs import TOWERS  # noqa: E402

DEV = "cuda"


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


ROWS = []


def row(name, ms_ours, ms_lib, lib, work=None, unit=""):
    r = {"op": name, "ms_b200": round(ms_ours, 4), "ms_torch": round(ms_lib, 4), "torch_op": lib,
         "speedup": round(ms_lib / ms_ours, 3)}
    if work:
        r["b200_" + unit] = round(work / ms_ours, 1)
        r["torch_" + unit] = round(work / ms_lib, 1)
    ROWS.append(r)
    print(json.dumps(r), flush=True)


def attention(B, T, nh, hd, tag):
    g = torch.Generator().manual_seed(0)
    hs = ops.attn_head_stride(hd)
    qkv = torch.zeros(B * T, 3, nh, hs, dtype=torch.bfloat16)
    qkv[..., :hd] = (torch.randn(B * T, 3, nh, hd, generator=g) * 1.5).bfloat16()
    qkv = qkv.view(B * T, 3 * nh * hs).to(DEV)
    out = torch.empty(B * T, nh * hd, dtype=torch.bfloat16, device=DEV)
    scale = hd ** -0.5
    ms = timeit(lambda: ops.attention_fwd(qkv, B, T, nh, hd, scale, out=out))
    # the library path as HF calls it: q/k/v [B, nh, T, hd] views of separate projections (transpose(1,2) views)
    x = qkv.view(B, T, 3, nh, hs)[..., :hd]
    q, k, v = [x[:, :, i].transpose(1, 2) for i in range(3)]
    if hs != hd:
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
    ms_lib = timeit(lambda: F.scaled_dot_product_attention(q, k, v, scale=scale))
    fl = 4.0 * B * nh * T * T * hd / 1e9
    row(f"attention {tag} (B{B} T{T} h{nh}x{hd})", ms, ms_lib, "F.scaled_dot_product_attention bf16", fl, "tflops")
    from torch.nn.attention import SDPBackend, sdpa_kernel
    for be in (SDPBackend.FLASH_ATTENTION, SDPBackend.CUDNN_ATTENTION, SDPBackend.EFFICIENT_ATTENTION):
        try:
            with sdpa_kernel(be):
                ms_be = timeit(lambda: F.scaled_dot_product_attention(q, k, v, scale=scale))
            row(f"attention {tag} vs {be.name}", ms, ms_be, be.name, fl, "tflops")
        except Exception as e:  # backend unavailable for this shape
            print(json.dumps({"op": f"attention {tag} vs {be.name}", "unavailable": str(e)[:120]}), flush=True)


def gemms(images, tower):
    c = TOWERS[tower]
    M, D, I = images * c.num_tokens, c.hidden_size, c.intermediate_size
    g = torch.Generator().manual_seed(0)
    r = lambda *s: (torch.randn(*s, generator=g) * 0.05).bfloat16().to(DEV)  # noqa: E731
    x, mlp = r(M, D), r(M, I)
    for name, a, N, K, act, res in (("qkv", x, 3 * D, D, None, False), ("out-proj +res", x, D, D, None, True),
                                    ("fc1 +act", x, I, D, c.hidden_act, False), ("fc2 +res", mlp, D, I, None, True)):
        w, b = r(N, K), r(N)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
        resid = r(M, N) if res else None
        ms = timeit(lambda: ops.gemm_bias_act(a, w, b, act=ops.ACT_BY_NAME[act], residual=resid, out=out))

        def lib():
            y = F.linear(a, w, b)
            if act == "quick_gelu":
                y = y * torch.sigmoid(1.702 * y)
            elif act == "gelu_pytorch_tanh":
                y = F.gelu(y, approximate="tanh")
            if res:
                y = resid + y
            return y
        ms_lib = timeit(lib)
        row(f"{tower} {name} (M{M} N{N} K{K})", ms, ms_lib, "F.linear (+act/+res kernels)", 2.0 * M * N * K / 1e9, "tflops")


def layernorm(images, tower):
    c = TOWERS[tower]
    M, D = images * c.num_tokens, c.hidden_size
    x = torch.randn(M, D, device=DEV).bfloat16()
    w, b = torch.randn(D, device=DEV).bfloat16(), torch.randn(D, device=DEV).bfloat16()
    out = torch.empty_like(x)
    ms = timeit(lambda: ops.layernorm(x, w, b, c.layer_norm_eps, out=out))
    ms_lib = timeit(lambda: F.layer_norm(x, (D,), w, b, c.layer_norm_eps))
    row(f"{tower} layernorm (M{M} D{D})", ms, ms_lib, "F.layer_norm", 4.0 * M * D / 1e6, "gbs")


def patch(images, tower):
    c = TOWERS[tower]
    D, p = c.hidden_size, c.patch_size
    px = torch.randn(images, 3, c.image_size, c.image_size, device=DEV)
    w = (torch.randn(D, 3, p, p, device=DEV) * 0.02).bfloat16()
    kpad = (c.patch_k + 63) // 64 * 64
    wp = torch.zeros(D, kpad, dtype=torch.bfloat16, device=DEV)
    wp[:, :c.patch_k] = w.reshape(D, -1)
    pos = torch.randn(c.num_tokens, D, device=DEV).bfloat16()
    cls = torch.randn(D, device=DEV).bfloat16() if c.has_cls else None
    bias = torch.randn(D, device=DEV).bfloat16() if c.conv_bias else None
    out = torch.empty(images, c.num_tokens, D, dtype=torch.bfloat16, device=DEV)
    ms = timeit(lambda: ops.patch_embed(px, wp, bias, pos, cls, p, out=out))

    def lib():   # HF:models/clip/modeling_clip.py:202-219
        y = F.conv2d(px.to(torch.bfloat16), w, bias, stride=p).flatten(2).transpose(1, 2)
        if cls is not None:
            y = torch.cat([cls.expand(images, 1, -1), y], dim=1)
        return y + pos.unsqueeze(0)
    ms_lib = timeit(lib)
    row(f"{tower} patch-embed (B{images})", ms, ms_lib, "cast + cuDNN conv2d + cat + add",
        2.0 * images * c.num_patches * D * c.patch_k / 1e9, "tflops")


def splice(rows, S, per_row_tokens, H, V, tag):
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, V - 4, (rows, S), generator=g)
    att = V - 2
    for r_ in range(rows):
        ids[r_, 1:1 + per_row_tokens] = att
    ids = ids.to(DEV)
    table = (torch.randn(V, H, device=DEV) * 0.02).bfloat16()
    n = rows * per_row_tokens
    proj = torch.randn(n, H, device=DEV).bfloat16()
    ms = timeit(lambda: ops.embed_gather(table, ids))
    ms_lib = timeit(lambda: F.embedding(ids, table))
    row(f"embedding {tag} ({rows}x{S}x{H})", ms, ms_lib, "F.embedding", 2.0 * rows * S * H * 2 / 1e6, "gbs")
    emb = ops.embed_gather(table, ids)
    dest, _, _ = ops.attach_positions(ids, att, expected_rows=n, want_src_of_pos=False)
    pos = (ids == att).nonzero()
    bi, tr = pos[:, 0].contiguous(), pos[:, 1].contiguous()
    ms = timeit(lambda: ops.splice_scatter(emb.view(-1, H), proj, dest))
    def lib():
        emb[bi, tr] = proj.view(-1, H).to(emb.dtype)     # model.py:441-444
    ms_lib = timeit(lib)
    row(f"splice {tag} ({n} rows x {H})", ms, ms_lib, "index_put_ (model.py:441-444)", 2.0 * n * H * 2 / 1e6, "gbs")
    ms = timeit(lambda: ops.attach_positions(ids, att, expected_rows=n, want_src_of_pos=False))
    ms_lib = timeit(lambda: (ids == att).nonzero())
    row(f"attach positions {tag}", ms, ms_lib, "(ids == att).nonzero()  [host sync inside]")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=256)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    attention(a.images, 577, 16, 64, "cfg2")
    attention(64, 729, 16, 72, "cfg3")
    attention(1280, 50, 12, 64, "cfg1/cfg5")
    gemms(a.images, "clip_l14_336")
    gemms(64, "siglip_so400m_384")
    gemms(1280, "clip_b32")
    layernorm(a.images, "clip_l14_336")
    patch(a.images, "clip_l14_336")
    patch(64, "siglip_so400m_384")
    patch(256, "clip_b32")
    splice(a.images, 642, 576, 4096, 128_258, "cfg2")
    splice(64, 8192, 2304, 4096, 128_258, "cfg4")
    print("\n%-52s %10s %10s %8s" % ("op", "b200 ms", "torch ms", "x"))
    for r in ROWS:
        print("%-52s %10.4f %10.4f %8.2f" % (r["op"][:52], r["ms_b200"], r["ms_torch"], r["speedup"]))


if __name__ == "__main__":
    main()
