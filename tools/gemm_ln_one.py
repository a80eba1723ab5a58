"""One LayerNorm-folded QKV GEMM (cfg2 shape) for an ncu capture:  python tools/gemm_ln_one.py [images] [fold|plain]"""
import sys, torch
sys.path.insert(0, ".")
from multimeditron_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
mode = sys.argv[2] if len(sys.argv) > 2 else "fold"
M, D = B * 577, 1024
g = torch.Generator().manual_seed(0)
r = lambda *s: (torch.randn(*s, generator=g) * 0.05).bfloat16().cuda()
N = 4 * D if (len(sys.argv) > 3 and sys.argv[3] == "fc1") else 3 * D
act = 1 if N == 4 * D else 0
x, w, b = r(M, D), r(N, D), r(N)
stats = torch.rand(M, ops.row_stats_slots(D), 2, device="cuda")
wf, cb = ops.fold_layernorm(w, b, 1 + r(D), r(D))
out = torch.empty(M, N, device="cuda").bfloat16()
for _ in range(3):
    if mode == "fold":
        ops.gemm_ln_bias_act(x, wf, cb, act=act, out=out, ln_stats=stats)
    else:
        ops.gemm_bias_act(x, w, b, act=act, out=out)
torch.cuda.synchronize()
