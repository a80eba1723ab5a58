"""Attention microbenchmark (GPU box): times mm_attention_fwd on the bench shapes with CUDA events and checks it
against torch SDPA (fp32) on a slice.  Usage: python tools/attn_bench.py [--shape cfg2|cfg3|cfg1] [--iters 20]"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multimeditron_b200 import ops  # noqa: E402

SHAPES = {  # B, T, heads, head_dim
    "cfg2": (256, 577, 16, 64),
    "cfg3": (128, 729, 16, 72),
    "cfg1": (512, 50, 12, 64),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="cfg2")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--T", type=int, default=0)
    a = ap.parse_args()
    B, T, nh, hd = SHAPES[a.shape]
    T = a.T or T
    dev = torch.device("cuda:0")
    hs = ops.attn_head_stride(hd) if hasattr(ops, "attn_head_stride") else hd
    g = torch.Generator().manual_seed(0)
    qkv = torch.zeros(B * T, 3, nh, hs, dtype=torch.bfloat16)
    qkv[..., :hd] = (torch.randn(B * T, 3, nh, hd, generator=g) * 1.5).bfloat16()
    qkv = qkv.view(B * T, 3 * nh * hs).to(dev)
    scale = hd ** -0.5
    out = torch.empty(B * T, nh * hd, dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        ops.attention_fwd(qkv, B, T, nh, hd, scale, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        ops.attention_fwd(qkv, B, T, nh, hd, scale, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    flops = 4.0 * B * nh * T * T * hd
    # correctness on the first and last 2 images
    err = 0.0
    for b0 in (0, B - 2):
        x = qkv.view(B, T, 3, nh, hs)[b0:b0 + 2, :, :, :, :hd].float()
        q, k, v = [x[:, :, i].transpose(1, 2) for i in range(3)]
        ref = F.scaled_dot_product_attention(q, k, v, scale=scale).transpose(1, 2).reshape(2 * T, nh * hd)
        got = out.view(B, T, nh * hd)[b0:b0 + 2].reshape(2 * T, nh * hd).float()
        err = max(err, (got - ref).abs().max().item() / ref.abs().max().item())
    print(json.dumps({"shape": a.shape, "B": B, "T": T, "heads": nh, "hd": hd, "ms": ms, "tflops": flops / ms / 1e9,
                      "rel_err": err}))


if __name__ == "__main__":
    main()
