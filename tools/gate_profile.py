"""Per-launch CUDA-event times of the native ResNet-50 gate (GPU box).  python tools/gate_profile.py [batch]"""
import sys, torch
sys.path.insert(0, ".")
from multimeditron_b200 import ops
from multimeditron_b200.gating import B200ResNet50Gate, ResNet50Gate
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
tv = ResNet50Gate(5, device="cpu", dtype=torch.float32, fold_bn=False)
g = B200ResNet50Gate({"resnet." + k: v for k, v in tv.resnet.state_dict().items()}, 5, device="cuda")
x = torch.randn(B, 3, 224, 224, device="cuda").bfloat16()
for _ in range(3):
    g(x)
torch.cuda.synchronize()
ops.profile_start()
g(x)
prof = ops._PROFILE
ops._PROFILE = None
torch.cuda.synchronize()
tot = 0.0
for name, evs in prof.items():
    ms = sum(a.elapsed_time(b) for a, b, _ in evs)
    tot += ms
    print(f"{name:22s} n={len(evs):3d} {ms:8.3f} ms")
    if name in ("mm_gemm_bias_act", "mm_conv_rows"):
        for a, b, w in evs:
            t = a.elapsed_time(b)
            if t > 0.10:
                print(f"      {t:7.3f} ms  work {w / 1e9:9.2f} G  -> {w / t / 1e9:8.1f} T/s|GB/s")
print("sum", tot)
