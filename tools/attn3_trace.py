"""Phase trace of the 3-stream attention kernel (GPU box).  Needs a library built with -DMM_ATTN_TRACE
(MM_TRACE_LIB=tools/ab/libmm_trace.so).  Prints, for block 0, clock64 deltas of warp q=0 of every softmax stream
(SMs: 0 s_full seen, 1 S in regs, 2 exps done, 3 pv_done seen, 4 P stored) and of every MMA issuer (6 = QK issued,
8 = PV issued)."""
import ctypes as C, os, sys
import torch
lib = C.CDLL(os.environ.get("MM_TRACE_LIB", "tools/ab/libmm_trace.so"))
T = int(sys.argv[1]) if len(sys.argv) > 1 else 577
N = int(sys.argv[2]) if len(sys.argv) > 2 else 300
B, nh, hd = 64, 16, 64
qkv = (torch.randn(B * T, 3 * nh * hd, device="cuda") * 1.5).bfloat16()
out = torch.empty(B * T, nh * hd, device="cuda", dtype=torch.bfloat16)
buf = torch.zeros(6 * 512 * 2, dtype=torch.int64, device="cuda")
lib.mm_attention_fwd.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.c_void_p]
lib.mm_attention3_trace_set.argtypes = [C.c_void_p]
for i in range(2):
    assert lib.mm_attention3_trace_set(buf.data_ptr() if i == 1 else None) == 0
    assert lib.mm_attention_fwd(qkv.data_ptr(), out.data_ptr(), B, T, nh, hd, hd ** -0.5, None) == 0
    torch.cuda.synchronize()
ev = buf.view(6, 512, 2).cpu()
rows = []
for role in range(6):
    for k in range(512):
        i, c = int(ev[role, k, 0]), int(ev[role, k, 1])
        if c:
            rows.append((c, role, i))
rows.sort()
t0 = rows[0][0]
names = {0: "s_full", 1: "S_regs", 2: "exps", 3: "pv_done", 4: "P_st", 6: "QK issued", 8: "PV issued"}
last = {}
for c, role, i in rows[:N]:
    who = f"SM{role}" if role < 3 else f"MMA{role - 3}"
    d = c - last.get(role, c)
    last[role] = c
    print(f"{c - t0:8d}  {'    ' * role}{who} {names[i]} (+{d})")
