"""Bring-up probe for the sm_100a kernels: small self-checking cases with diagnostics (run on the GPU box).

    timeout -k 5 120 python tools/gpu_probe.py splice|ln|gemm|patch|attn|tower
"""
import sys
import os
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multimeditron_b200 import ops  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)


def report(name, got, ref, tol=2e-2):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs()
    scale = ref.abs().max().item() + 1e-9
    bad = err > tol * scale + 1.6e-2 * ref.abs()
    ok = not bool(bad.any()) and bool(torch.isfinite(got).all())
    print(f"[{'OK ' if ok else 'BAD'}] {name}: maxerr {err.max().item():.4e} scale {scale:.3e} nbad {int(bad.sum())}/{bad.numel()}",
          flush=True)
    if not ok and got.dim() == 2:
        rows = bad.any(1).nonzero().flatten()
        cols = bad.any(0).nonzero().flatten()
        print("    bad rows (first 16):", rows[:16].tolist(), " count", rows.numel())
        print("    bad cols (first 16):", cols[:16].tolist(), " count", cols.numel())
        r, c = (rows[0].item(), cols[0].item()) if rows.numel() else (0, 0)
        print("    got[r, c:c+8]", got[r, c:c + 8].tolist())
        print("    ref[r, c:c+8]", ref[r, c:c + 8].tolist())
    return ok


def probe_splice():
    B, S, H, att = 4, 1000, 256, 7
    ids = torch.randint(8, 100, (B, S), device=dev)
    ids[torch.rand(B, S, device=dev) < 0.3] = att
    n = int((ids == att).sum())
    dest, sop, cnt = ops.attach_positions(ids, att, expected_rows=n)
    ref = (ids.view(-1) == att).nonzero().flatten().int()
    print("attach_positions:", torch.equal(dest, ref), int(cnt), n)
    table = torch.randn(100, H, device=dev).bfloat16()
    src = torch.randn(n, H, device=dev).bfloat16()
    emb = ops.embed_gather(table, ids)
    print("embed_gather:", torch.equal(emb, table[ids]))
    want = table[ids].clone().view(-1, H)
    want[ref.long()] = src
    ops.splice_scatter(emb.view(-1, H), src, dest)
    print("splice_scatter:", torch.equal(emb.view(-1, H), want))
    fused = ops.embed_splice_fused(table, ids, sop, src)
    print("embed_splice_fused:", torch.equal(fused.view(-1, H), want))


def probe_ln():
    for M, D in [(77, 768), (300, 1024), (50, 1152), (33, 128)]:
        x = torch.randn(M, D, device=dev).bfloat16()
        g = (1 + 0.1 * torch.randn(D, device=dev)).bfloat16()
        b = (0.1 * torch.randn(D, device=dev)).bfloat16()
        y = ops.layernorm(x, g, b, 1e-5)
        ref = torch.nn.functional.layer_norm(x.float(), (D,), g.float(), b.float(), 1e-5)
        report(f"layernorm {M}x{D}", y, ref)


def gemm_case(M, N, K, bias=False, act=0, residual=False, rowmap=False, name=""):
    a = torch.randn(M, K, device=dev).bfloat16()
    w = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bs = (torch.randn(N, device=dev) * 0.5).bfloat16() if bias else None
    res = torch.randn(M, N, device=dev).bfloat16() if residual else None
    ref = a.float() @ w.float().t()
    if bias:
        ref = ref + bs.float()
    if act == 1:
        ref = ref * torch.sigmoid(1.702 * ref)
    elif act == 2:
        ref = torch.nn.functional.gelu(ref)
    elif act == 3:
        ref = torch.nn.functional.gelu(ref, approximate="tanh")
    if residual:
        ref = ref + res.float()
    if rowmap:
        perm = torch.randperm(M, device=dev).int()
        perm[::7] = -1
        out = torch.zeros(M, N, device=dev).bfloat16()
        ops.gemm_bias_act(a, w, bs, act, residual=res, out=out, out_row_map=perm)
        want = torch.zeros(M, N, device=dev)
        keep = perm >= 0
        want[perm[keep].long()] = ref[keep]
        got = out
        ref = want
    else:
        got = ops.gemm_bias_act(a, w, bs, act, residual=res)
    torch.cuda.synchronize()
    return report(f"gemm {name} M{M} N{N} K{K} bias{int(bias)} act{act} res{int(residual)} map{int(rowmap)}", got, ref)


def probe_gemm():
    gemm_case(128, 256, 64, name="1tile-1k")
    gemm_case(128, 256, 256, name="1tile-4k")
    gemm_case(128, 256, 1024, name="1tile-16k(ring wrap)")
    gemm_case(256, 512, 512, name="4tiles")
    gemm_case(300, 264, 200, name="tails")
    gemm_case(1000, 768, 3072, bias=True, act=1, name="fc1-like")
    gemm_case(1000, 1024, 1024, bias=True, residual=True, name="outproj-like")
    gemm_case(577 * 3, 1024, 4096, bias=True, act=2, residual=True, rowmap=True, name="map")
    gemm_case(50, 1152, 4304, bias=True, act=3, name="siglip-odd")
    gemm_case(148 * 128 * 3 + 5, 512, 128, bias=True, name="many-tiles")  # multi-tile persistence + tmem double buffer
    # throughput sample
    M, N, K = 16384, 4096, 4096
    a = torch.randn(M, K, device=dev).bfloat16()
    w = torch.randn(N, K, device=dev).bfloat16()
    out = torch.empty(M, N, device=dev).bfloat16()
    for _ in range(3):
        ops.gemm_bias_act(a, w, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        ops.gemm_bias_act(a, w, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"gemm {M}x{N}x{K}: {ms:.3f} ms  {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
    e0.record()
    for _ in range(10):
        torch.matmul(a, w.t())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"torch.matmul same: {ms:.3f} ms  {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)


def probe_patch():
    import torch.nn.functional as F
    for (B, img, p, D, cls, bias, dt) in [(2, 32, 8, 128, True, False, torch.float32), (3, 224, 32, 768, True, False, torch.float32),
                                          (2, 336, 14, 1024, True, False, torch.float32), (2, 56, 14, 256, False, True, torch.bfloat16)]:
        G = img // p
        T = G * G + (1 if cls else 0)
        px = torch.randn(B, 3, img, img, device=dev).to(dt)
        w = (torch.randn(D, 3, p, p, device=dev) * 0.02).bfloat16()
        bs = (torch.randn(D, device=dev) * 0.1).bfloat16() if bias else None
        pos = (torch.randn(T, D, device=dev) * 0.1).bfloat16()
        c = (torch.randn(D, device=dev) * 0.1).bfloat16() if cls else None
        K = 3 * p * p
        kpad = (K + 63) // 64 * 64
        wp = torch.zeros(D, kpad, device=dev).bfloat16()
        wp[:, :K] = w.reshape(D, -1)
        out = ops.patch_embed(px, wp, bs, pos, c, p)
        ref = F.conv2d(px.bfloat16().float(), w.float(), bs.float() if bias else None, stride=p).flatten(2).transpose(1, 2)
        if cls:
            ref = torch.cat([c.float().expand(B, 1, D), ref], 1)
        ref = ref + pos.float()
        torch.cuda.synchronize()
        report(f"patch_embed B{B} img{img} p{p} D{D} cls{int(cls)}", out.reshape(B * T, D), ref.reshape(B * T, D))


def probe_attn():
    import torch.nn.functional as F
    for (B, T, nh) in [(1, 50, 2), (2, 128, 2), (2, 200, 3), (2, 577, 4), (40, 577, 16)]:
        D = nh * 64
        qkv = torch.randn(B * T, 3 * D, device=dev).bfloat16()
        out = ops.attention_fwd(qkv, B, T, nh, 64, 0.125)
        q, k, v = [t.view(B, T, nh, 64).transpose(1, 2).float() for t in qkv.view(B, T, 3 * D).split(D, -1)]
        ref = F.scaled_dot_product_attention(q, k, v, scale=0.125).transpose(1, 2).reshape(B * T, D)
        torch.cuda.synchronize()
        report(f"attention B{B} T{T} heads{nh}", out, ref, tol=1e-2)
    B, T, nh = 64, 577, 16
    D = nh * 64
    qkv = torch.randn(B * T, 3 * D, device=dev).bfloat16()
    out = torch.empty(B * T, D, device=dev).bfloat16()
    for _ in range(3):
        ops.attention_fwd(qkv, B, T, nh, 64, 0.125, out=out)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        ops.attention_fwd(qkv, B, T, nh, 64, 0.125, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"attention B{B} T{T} h{nh}: {ms:.3f} ms  {4 * B * nh * T * T * 64 / ms / 1e9:.1f} TFLOP/s", flush=True)


def probe_tower():
    from multimeditron_b200 import synth
    from multimeditron_b200.configs import TOWERS
    from multimeditron_b200.vit import TowerRunner, TowerWeights
    from oracle import vit_oracle
    for name, H, B in [("tiny", 64, 4), ("clip_b32", 512, 8)]:
        cfg = TOWERS[name]
        sd = synth.make_modality_state_dict(cfg, H, seed=5)
        px = synth.make_pixels(cfg, B, seed=5)
        ref = vit_oracle.image_modality_forward(sd, cfg, px)
        r = TowerRunner(TowerWeights(cfg, sd, dev))
        y = r.forward(px.to(dev))
        torch.cuda.synchronize()
        report(f"tower {name}", y.reshape(-1, H).cpu(), ref.reshape(-1, H), tol=3e-2)


def probe_gemmshapes():
    """Per-shape throughput of the GEMMs of one CLIP-L/14-336 step (M = 256*577)."""
    M = 256 * 577
    shapes = [("qkv", 3072, 1024, 0, False), ("out", 1024, 1024, 0, True), ("fc1", 4096, 1024, 1, False),
              ("fc2", 1024, 4096, 0, True), ("fc1-noact", 4096, 1024, 0, False), ("out-nores", 1024, 1024, 0, False),
              ("proj0", 1024, 1024, 2, False), ("proj2", 4096, 1024, 2, False), ("proj4", 4096, 4096, 0, False)]
    for name, N, K, act, res in shapes:
        a = torch.randn(M, K, device=dev).bfloat16()
        w = (torch.randn(N, K, device=dev) * 0.02).bfloat16()
        b = torch.randn(N, device=dev).bfloat16()
        r = torch.randn(M, N, device=dev).bfloat16() if res else None
        out = torch.empty(M, N, device=dev).bfloat16()
        for _ in range(2):
            ops.gemm_bias_act(a, w, b, act, residual=r, out=out)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(5):
            ops.gemm_bias_act(a, w, b, act, residual=r, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"{name:10s} M{M} N{N} K{K} act{act} res{int(res)}: {ms:.3f} ms {2 * M * N * K / ms / 1e9:.0f} TFLOP/s", flush=True)
        del a, w, r, out


if __name__ == "__main__":
    which = sys.argv[1]
    t = time.time()
    {"splice": probe_splice, "ln": probe_ln, "gemm": probe_gemm, "patch": probe_patch, "attn": probe_attn,
     "tower": probe_tower, "gemmshapes": probe_gemmshapes}[which]()
    torch.cuda.synchronize()
    print(f"-- {which} done in {time.time() - t:.1f}s", flush=True)
