"""One fused ResBlock unit launch per kernel size at HiFi-GAN's 32-channel (default) or 64-channel stage shape (B = 32,
163 840 / 81 920 rows), timed next to the two launches it replaces; with --ncu only the fused launches (for
`ncu -k regex:resunit`).  usage: python scripts/resunit_one.py [32|64] [--ncu]"""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
C = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 32
B, S = 32, 640 * 8192 // C
L = ops.ACT_LRELU
y = torch.randn(B, S, C, device="cuda:0")
xp = ops.split_bf16(y, act=L, slope=0.1)
tp, yp, out = torch.empty_like(xp), torch.empty_like(xp), torch.empty_like(y)


def timed(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


for k, dil in (((3, 1), (7, 3), (11, 5)) if C == 32 else ((3, 1), (3, 5))):
    W1 = ops.pack_conv(torch.randn(C, C, k, device="cuda:0") / math.sqrt(C * k))
    W2 = ops.pack_conv(torch.randn(C, C, k, device="cuda:0") / math.sqrt(C * k))
    b1, b2 = torch.randn(C, device="cuda:0"), torch.randn(C, device="cuda:0")

    def fused():
        ops.resunit_x3(xp, y, W1, b1, W2, b2, dilation=dil, slope=0.1, out=out, act2=L, slope2=0.1, out_pair=yp)

    def two():
        ops.conv1d_x3(xp, W1, b1, dilation=dil, pad=(k - 1) // 2 * dil, want_y=False, act2=L, slope2=0.1, out_pair=tp)
        ops.conv1d_x3(tp, W2, b2, pad=(k - 1) // 2, residual=y, out=out, act2=L, slope2=0.1, out_pair=yp)

    if "--ncu" in sys.argv:
        for _ in range(3):
            fused()
        torch.cuda.synchronize()
    else:
        tf, t2 = timed(fused), timed(two)
        gb = B * S * C * 16 / 1e9
        print(f"k={k} dil={dil}: fused {tf:7.1f} us ({gb / tf * 1e3:5.2f} TB/s on 16 B/element)   two launches {t2:7.1f} us", flush=True)
