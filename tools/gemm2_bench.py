"""CTA-pair (cta_group::2) GEMM experiment: correctness vs the 1-CTA kernel and per-shape timing (cfg2 shapes).
    timeout 120 python tools/gemm2_bench.py [images]"""
import ctypes, sys, torch
sys.path.insert(0, ".")
from multimeditron_b200 import ops, _lib

lib = _lib.load()
f = lib.mm_gemm2_bias
f.restype = ctypes.c_int
f.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
              ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]


def gemm2(a, w, bias, out):
    rc = f(a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), bias.data_ptr() if bias is not None else None,
           out.data_ptr(), out.stride(0), a.shape[0], w.shape[0], a.shape[1],
           torch.cuda.current_stream().cuda_stream)
    if rc:
        raise RuntimeError(f"mm_gemm2_bias rc={rc}")
    return out


dev = "cuda"
g = torch.Generator().manual_seed(0)
r = lambda *s: (torch.randn(*s, generator=g) * 0.05).bfloat16().to(dev)

# ---- correctness: ragged and aligned shapes, bit-exact vs the 1-CTA kernel
for (M, N, K) in [(256, 256, 64), (256, 256, 512), (512, 512, 1024), (300, 264, 200), (1000, 1024, 1024), (577 * 4, 3072, 1024)]:
    a, w, b = r(M, K), r(N, K), r(N)
    want = ops.gemm_bias_act(a, w, b)
    got = torch.full((M, N), 7.0, device=dev).bfloat16()
    gemm2(a, w, b, got)
    torch.cuda.synchronize()
    ok = torch.equal(want, got)
    print(f"check M={M} N={N} K={K}: {'bit-exact' if ok else 'MISMATCH max ' + str((want.float() - got.float()).abs().max().item())}", flush=True)
    if not ok:
        sys.exit(1)

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
M, D, I = B * 577, 1024, 4096
x, mlp = r(M, D), r(M, I)
wqkv, wo, w1, w2 = r(3 * D, D), r(D, D), r(I, D), r(D, I)
bq, bo, b1, b2 = r(3 * D), r(D), r(I), r(D)
qkv, out, h = (torch.empty(M, n, device=dev).bfloat16() for n in (3 * D, D, I))


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
    print(f"{name:26s} {ms:7.3f} ms  {flop / ms / 1e9:7.1f} TF/s", flush=True)


for name, a, w, b, o in [("qkv", x, wqkv, bq, qkv), ("out", x, wo, bo, out), ("fc1", x, w1, b1, h), ("fc2", mlp, w2, b2, out)]:
    flop = 2.0 * a.shape[0] * w.shape[0] * a.shape[1]
    t(f"{name} 1-CTA", lambda: ops.gemm_bias_act(a, w, b, out=o), flop)
    t(f"{name} CTA-pair", lambda: gemm2(a, w, b, o), flop)
    t(f"{name} cuBLAS", lambda: torch.addmm(b, a, w.t(), out=o), flop)
