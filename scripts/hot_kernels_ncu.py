"""The kernels that carry the cfg2 step, one launch each at their cfg2 shapes (for one ncu --set full pass):
conv1d_tf32_ts2_kernel (CTA pair; FFN 256->1024 k=9, B=64 S=696), conv1d_tf32_ts_kernel<128,4> (QKV 256->768 k=1),
attention_tc_kernel (8x32 heads, S=800), attention_tc128_kernel (2x128 heads, S=696)."""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
B, S, Cin, Cout, k = 64, 696, 256, 1024, 9
x = torch.randn(B, S, Cin, device=DEV)
W = ops.pack_conv(torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k))
bias = torch.randn(Cout, device=DEV)
lens = torch.randint(500, S + 1, (B,), dtype=torch.int32, device=DEV)
q32 = torch.randn(B, 800, 3 * 256, device=DEV)
l32 = torch.full((B,), 800, dtype=torch.int32, device=DEV)
q128 = torch.randn(B, S, 3 * 256, device=DEV)
Wq = ops.pack_conv(torch.randn(768, Cin, 1, device=DEV) / math.sqrt(Cin))
bq = torch.randn(768, device=DEV)
for _ in range(2):
    ops.conv1d(x, W, bias, pad=k // 2, act=ops.ACT_RELU, lens=lens)
    ops.conv1d(x, Wq, bq)
    ops.attention(q32, l32, 8, backend="tc")
    ops.attention(q128, lens, 2, backend="tc")
torch.cuda.synchronize()
