"""Times the split-bf16 (bf16x3) convolution on HiFi-GAN's MRF shapes (dense B = 32, T = 640 frames) next to the 3xTF32
kernels of the fp32 mode.  usage: python scripts/x3_time.py [--few]"""
import math
import sys

import torch

sys.path.insert(0, ".")
from speakingstyle_b200 import ops  # noqa: E402

DEV = "cuda:0"
B, T = 32, 640
L = ops.ACT_LRELU
shapes = [(256, 8), (128, 64), (64, 128), (32, 256)]
cases = [(3, 1), (7, 1), (11, 1), (11, 5), (3, 5)] if "--few" not in sys.argv else [(3, 1), (11, 1)]


def run(fn, flops, tag):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    n = 5
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n
    print(f"{tag:60s} {ms * 1e3:9.1f} us  {flops / ms / 1e9:7.1f} TFLOP/s", flush=True)


for C, up in shapes:
    S = T * up
    x = torch.randn(B, S, C, device=DEV)
    xp = ops.split_bf16(x)
    tp = torch.empty_like(xp)
    yp = torch.empty_like(xp)
    y = torch.empty_like(x)
    xs = torch.zeros_like(x)
    for k, dil in cases:
        W = ops.pack_conv(torch.randn(C, C, k, device=DEV) / math.sqrt(C * k))
        bias = torch.randn(C, device=DEV)
        flops = 2.0 * B * S * C * C * k
        pad = (k * dil - dil) // 2
        run(lambda: ops.conv1d_x3(xp, W, bias, dilation=dil, pad=pad, want_y=False, act2=L, slope2=0.1, out_pair=tp), flops,
            f"x3 c1 C={C} S={S} k={k} dil={dil}")
        if dil == 1:
            run(lambda: ops.conv1d_x3(tp, W, bias, pad=pad, residual=x, out=y, act2=L, slope2=0.1, out_pair=yp), flops,
                f"x3 c2+res C={C} S={S} k={k}")
            run(lambda: ops.conv1d_x3(tp, W, bias, pad=pad, residual=x, want_y=False, accum=xs, accum_scale=1.0 / 3,
                                      accum_init=False), flops, f"x3 c2+res+mean C={C} S={S} k={k}")
    del x, xp, tp, yp, y, xs
    torch.cuda.empty_cache()
torch.cuda.synchronize()
