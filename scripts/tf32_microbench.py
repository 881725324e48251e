"""One 3xTF32 convolution launch at the ReferenceEncoder's biggest shape (for ncu)."""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
B, S, Cin, Cout, k = 64, 800, 1024, 1024, 3
x = torch.randn(B, S, Cin, device=DEV)
W = ops.pack_conv(torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k))
bias = torch.randn(Cout, device=DEV)
variants = sys.argv[1:] or ["ss", "ts"]
for var in variants:
    be = "tf32x3_" + var
    for _ in range(2): ops.conv1d(x, W, bias, pad=1, act=ops.ACT_RELU, backend=be)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); ops.conv1d(x, W, bias, pad=1, act=ops.ACT_RELU, backend=be); b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    print(f"{be} conv 1024->1024 k3 B=64 S=800: {ms:.3f} ms  {2.0*B*S*Cout*k*Cin/ms/1e9:.1f} TFLOP/s (algorithmic)")
