"""Patch-embed timing: staged (patch rows + CTA-pair GEMM) vs in-kernel gather.  python tools/patch_bench.py [images]"""
import sys, torch
sys.path.insert(0, ".")
from multimeditron_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
for img, p, D, cls in [(336, 14, 1024, True), (384, 14, 1152, False), (224, 32, 768, True)]:
    G = img // p
    T = G * G + (1 if cls else 0)
    px = torch.randn(B, 3, img, img, device="cuda")
    K = 3 * p * p
    wp = torch.zeros(D, (K + 63) // 64 * 64, device="cuda").bfloat16()
    wp[:, :K] = (torch.randn(D, K, device="cuda") * 0.02).bfloat16()
    pos = torch.randn(T, D, device="cuda").bfloat16()
    c = torch.randn(D, device="cuda").bfloat16() if cls else None
    out = torch.empty(B, T, D, device="cuda", dtype=torch.bfloat16)
    for staged in (False, True):
        for _ in range(3):
            ops.patch_embed(px, wp, None, pos, c, p, out=out, staged=staged)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.patch_embed(px, wp, None, pos, c, p, out=out, staged=staged)
        e1.record()
        torch.cuda.synchronize()
        print(f"img {img} patch {p} D {D} B {B}  {'staged' if staged else 'gather'}: {e0.elapsed_time(e1) / 10:.3f} ms", flush=True)
