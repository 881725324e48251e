"""Back-to-back timing of the short-reduction (k = 1) 3xTF32 convs at several batch sizes: separates a fixed
per-launch cost from the per-tile cost."""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
for (Cin, Cout, k) in ((256, 256, 1), (256, 768, 1), (1024, 256, 1), (256, 1024, 9)):
    W = ops.pack_conv(torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k))
    bias = torch.randn(Cout, device=DEV)
    for B in (2, 8, 16, 32, 64, 128):
        S = 696
        x = torch.randn(B, S, Cin, device=DEV)
        y = torch.empty(B, S, Cout, device=DEV)
        for _ in range(3): ops.conv1d(x, W, bias, pad=k // 2, out=y)
        torch.cuda.synchronize()
        n = 20
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n): ops.conv1d(x, W, bias, pad=k // 2, out=y)
        b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) / n * 1e3
        tiles = math.ceil(B * S / 128) * math.ceil(Cout / 128)
        print(f"{Cin:5d}->{Cout:4d} k={k} B={B:3d}: {us:8.1f} us  tiles={tiles:5d} rounds={math.ceil(tiles / 148):3d}  "
              f"{2.0 * B * S * Cin * Cout * k / us / 1e6:7.1f} TFLOP/s  us/round={us / math.ceil(tiles / 148):6.1f}")

print("--- rotating buffers (working set >> L2), residual read + row mask as in the FFT block ---")
for (Cin, Cout, k) in ((256, 256, 1), (256, 768, 1), (1024, 256, 1)):
    W = ops.pack_conv(torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k))
    bias = torch.randn(Cout, device=DEV)
    B, S, R = 64, 696, 6
    xs = [torch.randn(B, S, Cin, device=DEV) for _ in range(R)]
    rs = [torch.randn(B, S, Cout, device=DEV) for _ in range(R)]
    ys = [torch.empty(B, S, Cout, device=DEV) for _ in range(R)]
    for with_res in (False, True):
        for i in range(R): ops.conv1d(xs[i], W, bias, out=ys[i], residual=rs[i] if with_res else None)
        torch.cuda.synchronize()
        n = 24
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(n): ops.conv1d(xs[i % R], W, bias, out=ys[i % R], residual=rs[i % R] if with_res else None)
        b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) / n * 1e3
        mb = B * S * (Cin + Cout * (2 if with_res else 1)) * 4 / 1e6
        print(f"{Cin:5d}->{Cout:4d} k={k} residual={with_res}: {us:8.1f} us  {2.0 * B * S * Cin * Cout * k / us / 1e6:7.1f} TFLOP/s  "
              f"HBM {mb:.0f} MB -> {mb / us / 1e3:.2f} TB/s")
