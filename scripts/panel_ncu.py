"""One launch of the narrow-layer 3xTF32 conv at a HiFi-GAN stage shape: python scripts/panel_ncu.py C k [dil] [variant]"""
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
for _ in range(2):
    ops.conv1d(x, W, bias, k=k, dilation=dil, pad=(k - 1) * dil // 2, in_act=ops.ACT_LRELU, in_slope=0.1, out=t,
               backend="tf32x3_" + variant)
torch.cuda.synchronize()
