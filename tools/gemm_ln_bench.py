"""Per-shape timing of the encoder GEMMs: plain vs stats-emitting vs LayerNorm-folded epilogues (cfg2 shapes).
    python tools/gemm_ln_bench.py [images]        (set MM_B200_LIB=/path/to/other/libmm_b200.so to A/B a build;
    MM_B200_GEMM_PAIR=0 times the 1-CTA kernel instead of the CTA-pair kernel)"""
import sys, torch
sys.path.insert(0, ".")
from multimeditron_b200 import ops, _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
M, D, I = B * 577, 1024, 4096
dev = "cuda"
g = torch.Generator().manual_seed(0)
r = lambda *s: (torch.randn(*s, generator=g) * 0.05).bfloat16().to(dev)
x, attn, mlp = r(M, D), r(M, D), r(M, I)
wqkv, wo, w1, w2 = r(3 * D, D), r(D, D), r(I, D), r(D, I)
bq, bo, b1, b2 = r(3 * D), r(D), r(I), r(D)
gamma, beta = (1 + r(D)), r(D)
has_ln = hasattr(_lib.load(), "mm_gemm_ln_bias_act")
if not has_ln:   # old build under MM_B200_LIB: bind only what it exports
    pass
if has_ln:
    stats = torch.zeros(M, ops.row_stats_slots(D), 2, device=dev)
    wqkv_f, cbq = ops.fold_layernorm(wqkv, bq, gamma, beta)
    w1_f, cb1 = ops.fold_layernorm(w1, b1, gamma, beta)
qkv, out, h = torch.empty(M, 3 * D, device=dev).bfloat16(), torch.empty(M, D, device=dev).bfloat16(), torch.empty(M, I, device=dev).bfloat16()


def t(name, fn, flop, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name:34s} {ms:7.3f} ms  {flop / ms / 1e9:7.1f} TF/s", flush=True)


cases = [
    ("qkv   plain", lambda: ops.gemm_bias_act(x, wqkv, bq, out=qkv), 2.0 * M * 3 * D * D),
    ("out   plain +res", lambda: ops.gemm_bias_act(attn, wo, bo, residual=x, out=out), 2.0 * M * D * D),
    ("fc1   plain quickgelu", lambda: ops.gemm_bias_act(x, w1, b1, act=1, out=h), 2.0 * M * I * D),
    ("fc2   plain +res", lambda: ops.gemm_bias_act(mlp, w2, b2, residual=x, out=out), 2.0 * M * I * D),
]
if has_ln:
    cases += [
        ("qkv   ln-fold", lambda: ops.gemm_ln_bias_act(x, wqkv_f, cbq, out=qkv, ln_stats=stats), 2.0 * M * 3 * D * D),
        ("out   +res +stats", lambda: ops.gemm_ln_bias_act(attn, wo, bo, residual=x, out=out, stats_out=stats), 2.0 * M * D * D),
        ("fc1   ln-fold quickgelu", lambda: ops.gemm_ln_bias_act(x, w1_f, cb1, act=1, out=h, ln_stats=stats), 2.0 * M * I * D),
        ("fc2   +res +stats", lambda: ops.gemm_ln_bias_act(mlp, w2, b2, residual=x, out=out, stats_out=stats), 2.0 * M * I * D),
    ]
if has_ln:
    stats.copy_(torch.rand_like(stats) * 64 + 0.5)   # sane statistics (var > 0)
    cases += [
        ("fc1   plain no-act", lambda: ops.gemm_bias_act(x, w1, b1, act=0, out=h), 2.0 * M * I * D),
        ("fc1   ln-fold no-act", lambda: ops.gemm_ln_bias_act(x, w1_f, cb1, act=0, out=h, ln_stats=stats), 2.0 * M * I * D),
        ("qkv   plain quickgelu", lambda: ops.gemm_bias_act(x, wqkv, bq, act=1, out=qkv), 2.0 * M * 3 * D * D),
        ("qkv   ln-fold quickgelu", lambda: ops.gemm_ln_bias_act(x, wqkv_f, cbq, act=1, out=qkv, ln_stats=stats), 2.0 * M * 3 * D * D),
    ]
for _ in range(2):
    for c in cases:
        t(*c)
