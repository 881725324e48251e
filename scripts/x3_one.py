"""The four dilation-1 launches one HiFi-GAN ResBlock issues on the split-bf16 kernel at a stage shape (for ncu): c1,
c2 + residual (twice), c2 + residual + stage mean.  python scripts/x3_one.py C k [iters]"""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
C, k = int(sys.argv[1]), int(sys.argv[2])
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
B, T = 32, 640
S = T * (256 * 32 // C)
L = ops.ACT_LRELU
x = torch.randn(B, S, C, device="cuda:0")
xp = ops.split_bf16(x)
tp, yp, y, xs = torch.empty_like(xp), torch.empty_like(xp), torch.empty_like(x), torch.zeros_like(x)
W = ops.pack_conv(torch.randn(C, C, k, device="cuda:0") / math.sqrt(C * k))
bias = torch.randn(C, device="cuda:0")
pad = (k - 1) // 2
for _ in range(iters):
    ops.conv1d_x3(xp, W, bias, pad=pad, want_y=False, act2=L, slope2=0.1, out_pair=tp)
    ops.conv1d_x3(tp, W, bias, pad=pad, residual=x, out=y, act2=L, slope2=0.1, out_pair=yp)
    ops.conv1d_x3(tp, W, bias, pad=pad, residual=x, out=y, act2=L, slope2=0.1, out_pair=yp)
    ops.conv1d_x3(tp, W, bias, pad=pad, residual=x, want_y=False, accum=xs, accum_scale=1.0 / 3, accum_init=False)
torch.cuda.synchronize()
