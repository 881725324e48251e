"""One launch of a 3xTF32 conv at a given shape (for ncu): python scripts/conv_ncu.py Cin Cout k [residual]"""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
Cin, Cout, k = (int(v) for v in sys.argv[1:4])
res = len(sys.argv) > 4
B, S = 64, 696
x = torch.randn(B, S, Cin, device="cuda:0")
W = ops.pack_conv(torch.randn(Cout, Cin, k, device="cuda:0") / math.sqrt(Cin * k))
bias = torch.randn(Cout, device="cuda:0")
r = torch.randn(B, S, Cout, device="cuda:0") if res else None
for _ in range(3):
    ops.conv1d(x, W, bias, pad=k // 2, residual=r)
torch.cuda.synchronize()
