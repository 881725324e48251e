"""Timing of the narrow-layer conv at a HiFi-GAN stage shape: python scripts/panel_time.py C k [dil] [variant]"""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
C, k = int(sys.argv[1]), int(sys.argv[2])
dil = int(sys.argv[3]) if len(sys.argv) > 3 else 1
variant = sys.argv[4] if len(sys.argv) > 4 else "panel"
B, T = 32, 640
S = T * (256 * 32 // C)
x = torch.randn(B, S, C, device="cuda:0")
W = ops.pack_conv(torch.randn(C, C, k, device="cuda:0") / math.sqrt(C * k))
bias = torch.randn(C, device="cuda:0")
t = torch.empty_like(x)
f = lambda: ops.conv1d(x, W, bias, k=k, dilation=dil, pad=(k - 1) * dil // 2, in_act=ops.ACT_LRELU, in_slope=0.1, out=t, backend="tf32x3_" + variant)
for _ in range(3): f()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10): f()
b.record(); torch.cuda.synchronize()
us = a.elapsed_time(b) * 100
tiles = B * S / 128 / 148
print(f"C={C} k={k} dil={dil} {variant}: {us:.1f} us, {us * 1.9e3 / tiles:.0f} cycles/tile @1.9GHz, {2.0*B*S*C*C*k/us/1e6:.1f} TFLOP/s")
