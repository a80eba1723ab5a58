"""ResNet-50 gate (torchvision / cuDNN) timing variants on the GPU box: memory format, cudnn.benchmark, CUDA graph.
    python tools/gate_bench.py [batch]"""
import sys, torch
sys.path.insert(0, ".")
from multimeditron_b200.gating import ResNet50Gate
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = "cuda"
x = torch.randn(B, 3, 224, 224, device=dev).bfloat16()


def t(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for bench in (False, True):
    torch.backends.cudnn.benchmark = bench
    for fold in (False, True):
        g = ResNet50Gate(5, device=dev, fold_bn=fold)
        print(f"benchmark={bench} fold_bn={fold} channels_last bf16: {t(lambda: g(x)):.2f} ms", flush=True)
        g2 = ResNet50Gate(5, device=dev, fold_bn=fold)
        g2.resnet = g2.resnet.to(memory_format=torch.contiguous_format)
        xc = x.contiguous()
        f = lambda: torch.softmax(g2.resnet(xc), -1)
        print(f"benchmark={bench} fold_bn={fold} NCHW bf16: {t(f):.2f} ms", flush=True)
g = ResNet50Gate(5, device=dev, fold_bn=True)
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        g(x)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=s):
        out = g(x)
torch.cuda.current_stream().wait_stream(s)
print(f"CUDA graph, fold_bn, channels_last bf16: {t(gr.replay):.2f} ms")
