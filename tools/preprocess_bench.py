"""Throughput of mm_preprocess_images (device resize/crop/normalise) on device-resident raw images.
    python tools/preprocess_bench.py"""
import sys, torch
sys.path.insert(0, ".")
from multimeditron_b200 import preprocess as pre

dev = torch.device("cuda")
for name, B, H, W, S in [("336 identity", 256, 336, 336, 336), ("640x480 -> 336", 256, 480, 640, 336),
                         ("1024x1024 -> 224", 256, 1024, 1024, 224), ("2000x1500 -> 336", 64, 1500, 2000, 336)]:
    raw = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev)
    p = pre.DevicePreprocessor(S)
    for _ in range(3):
        out = p(raw, dev, torch.bfloat16)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        out = p(raw, dev, torch.bfloat16)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    # algorithmic bytes: source pixels inside the crop window's footprint read once + output written once
    oh, ow = pre.resize_target(H, W, S, False)
    frac = (S / oh) * (S / ow)
    bytes_alg = B * (H * W * 3 * frac + 3 * S * S * 2)
    print(f"{name:22s} B={B:4d}  {ms:7.3f} ms  {B / ms * 1e3:9.0f} img/s  {bytes_alg / ms / 1e6:7.1f} GB/s algorithmic "
          f"(source bytes in the crop footprint + bf16 output)", flush=True)
