"""Phase trace of the attention kernel (GPU box).  Builds csrc with -DMM_ATTN_TRACE into /tmp, runs one launch and
prints, for block 0, the clock64 deltas between phases of one softmax warp per stream and of the MMA issuer.
ids: softmax s*16 + {0 s_full seen, 1 S in regs, 2 exps done, 3 pv_done seen, 4 P stored, 5 p_full arrived};
MMA 6+s = S issued, 8+s = PV issued."""
import ctypes as C, os, subprocess, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = os.path.join(ROOT, "multimeditron_b200", "csrc")
so = os.environ.get("MM_TRACE_LIB", "/tmp/libmm_trace.so")
files = ["lib.cu", "splice.cu", "layernorm.cu", "gemm_sm100.cu", "attention_sm100.cu", "moe_fuse.cu", "preprocess.cu", "train_ops.cu"]
if not os.environ.get("MM_TRACE_LIB"):
  subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-DMM_ATTN_TRACE",
                       "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-shared", "-o", so, "-cudart", "static"]
                      + [os.path.join(src, f) for f in files])
lib = C.CDLL(so)
T = int(sys.argv[1]) if len(sys.argv) > 1 else 577
B, nh, hd = 64, 16, 64
qkv = (torch.randn(B * T, 3 * nh * hd, device="cuda") * 1.5).bfloat16()
out = torch.empty(B * T, nh * hd, device="cuda", dtype=torch.bfloat16)
buf = torch.zeros(3 * 1024 * 2, dtype=torch.int64, device="cuda")
lib.mm_attention_fwd.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.c_void_p]
lib.mm_attention_trace_set.argtypes = [C.c_void_p]
for i in range(2):
    assert lib.mm_attention_trace_set(buf.data_ptr() if i == 1 else None) == 0
    assert lib.mm_attention_fwd(qkv.data_ptr(), out.data_ptr(), B, T, nh, hd, hd ** -0.5, None) == 0
    torch.cuda.synchronize()
ev = buf.view(-1, 2).cpu().tolist()
ev = [(i, c) for i, c in ev if c != 0]
t0 = min(c for _, c in ev)
ev.sort(key=lambda e: e[1])
names = {0: "s_full", 1: "S_regs", 2: "exps", 3: "pv_done", 4: "P_st", 5: "p_full", 14: "max_done", 15: "token"}
ev = [e for e in ev if 6 <= e[0] <= 13] if len(sys.argv) > 3 else ev
for i, c in ev[:int(sys.argv[2]) if len(sys.argv) > 2 else 160]:
    if i in (6, 7):
        tag = f"MMA  S{i - 6} issued"
    elif i in (8, 9):
        tag = f"MMA  PV{i - 8} issued"
    elif i in (10, 11):
        tag = f"MMA  S{i - 10} begin"
    elif i in (12, 13):
        tag = f"MMA  PV{i - 12} begin"
    else:
        tag = f"SM{i // 16}   {names[i % 16]}"
    print(f"{c - t0:8d}  {tag}")
