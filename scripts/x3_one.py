"""One c1-style and one c2-style launch of the split-bf16 conv at a HiFi-GAN stage shape (for ncu):
python scripts/x3_one.py C k [dil]"""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
C, k = int(sys.argv[1]), int(sys.argv[2])
dil = int(sys.argv[3]) if len(sys.argv) > 3 else 1
B, T = 32, 640
S = T * (256 * 32 // C)
x = torch.randn(B, S, C, device="cuda:0")
xp = ops.split_bf16(x)
tp, yp, y = torch.empty_like(xp), torch.empty_like(xp), torch.empty_like(x)
W = ops.pack_conv(torch.randn(C, C, k, device="cuda:0") / math.sqrt(C * k))
bias = torch.randn(C, device="cuda:0")
pad = (k - 1) * dil // 2
for _ in range(2):
    ops.conv1d_x3(xp, W, bias, dilation=dil, pad=pad, want_y=False, act2=ops.ACT_LRELU, slope2=0.1, out_pair=tp)
    ops.conv1d_x3(tp, W, bias, pad=(k - 1) // 2, residual=x, out=y, act2=ops.ACT_LRELU, slope2=0.1, out_pair=yp)
torch.cuda.synchronize()
