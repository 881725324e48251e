"""First light for the CTA-pair conv kernel: parity against the single-CTA kernel (bit-exact expected: same
summation order) and timing."""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
torch.manual_seed(0)
for (B, S, Cin, Cout, k) in ((2, 300, 64, 128, 3), (3, 130, 256, 256, 1), (5, 50, 64, 128, 9), (64, 696, 256, 1024, 9), (64, 800, 1024, 1024, 3), (64, 696, 256, 256, 1)):
    x = torch.randn(B, S, Cin, device=DEV)
    W = ops.pack_conv(torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k))
    bias = torch.randn(Cout, device=DEV)
    res = torch.randn(B, S, Cout, device=DEV)
    lens = torch.randint(1, S + 1, (B,), dtype=torch.int32, device=DEV)
    y1 = ops.conv1d(x, W, bias, pad=k // 2, act=ops.ACT_RELU, residual=res, lens=lens, backend="tf32x3_ts")
    y2 = ops.conv1d(x, W, bias, pad=k // 2, act=ops.ACT_RELU, residual=res, lens=lens, backend="tf32x3_ts2")
    torch.cuda.synchronize()
    d = (y1 - y2).abs().max().item()
    ts = {}
    for be in ("tf32x3_ts", "tf32x3_ts2"):
        for _ in range(2): ops.conv1d(x, W, bias, pad=k // 2, backend=be)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5): ops.conv1d(x, W, bias, pad=k // 2, backend=be)
        b.record(); torch.cuda.synchronize()
        ts[be] = a.elapsed_time(b) / 5
    fl = 2.0 * B * S * Cin * Cout * k
    print(f"B={B} S={S} {Cin}->{Cout} k={k}: max|ts-ts2|={d:.2e}  ts {ts['tf32x3_ts']*1e3:8.1f} us ({fl/ts['tf32x3_ts']/1e9:6.1f} TF)  ts2 {ts['tf32x3_ts2']*1e3:8.1f} us ({fl/ts['tf32x3_ts2']/1e9:6.1f} TF)", flush=True)
