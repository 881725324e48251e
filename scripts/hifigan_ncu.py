"""One launch of each HiFi-GAN MRF convolution shape (for one `ncu --set full` pass, or timed with `--time`).

Stages of the V1 generator at a cfg3-like batch (B = 32, T = 640 mel frames): 256 ch x 8T, 128 ch x 64T, 64 ch x 128T,
32 ch x 256T rows; per stage the k = 3 / 7 / 11 ResBlocks, `c1` (leaky-ReLU on load, dilation 1 / 5) and `c2`
(dilation 1, + residual), in the fp32 mode (3xTF32 kernels) and the bf16 mode (conv1d_tc).
usage: python scripts/hifigan_ncu.py [fp32|bf16|both] [--time] [--few]"""
import math
import sys

import torch

sys.path.insert(0, ".")
from speakingstyle_b200 import ops  # noqa: E402

DEV = "cuda:0"
mode = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "both"
timed = "--time" in sys.argv
B, T = 32, 640
L = ops.ACT_LRELU
shapes = [(256, 8), (128, 64), (64, 128), (32, 256)]
cases = [(3, 1), (7, 1), (11, 1), (11, 5), (3, 5)] if "--few" not in sys.argv else [(3, 1), (11, 1)]


def run(fn, flops, tag):
    if not timed:
        fn()
        return
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
    t = torch.empty_like(x)
    y = torch.empty_like(x)
    xs = torch.zeros_like(x)
    xb = x.to(torch.bfloat16)
    for k, dil in cases:
        W = ops.pack_conv(torch.randn(C, C, k, device=DEV) / math.sqrt(C * k))
        bias = torch.randn(C, device=DEV)
        flops = 2.0 * B * S * C * C * k
        pad = (k * dil - dil) // 2
        if mode in ("fp32", "both"):
            # as hifigan.py issues them: c1 = leaky-ReLU on load and in the epilogue, c2 = plain input + residual (+ mean)
            run(lambda: ops.conv1d(x, W, bias, k=k, dilation=dil, pad=pad, in_act=L, in_slope=0.1, act=L, slope=0.1, out=t),
                flops, f"fp32 c1 C={C} S={S} k={k} dil={dil}")
            if dil == 1:
                run(lambda: ops.conv1d(t, W, bias, k=k, pad=pad, residual=x, out=y), flops,
                    f"fp32 c2+res C={C} S={S} k={k}")
                run(lambda: ops.conv1d(t, W, bias, k=k, pad=pad, residual=x, write_y=False,
                                       accum=xs, accum_scale=1.0 / 3, accum_init=False), flops,
                    f"fp32 c2+res+mean C={C} S={S} k={k}")
        if mode in ("bf16", "both"):
            run(lambda: ops.conv1d_tc(xb, W, bias, dilation=dil, pad=pad, want_y=False, act2=L, slope2=0.1), flops,
                f"bf16 c1 C={C} S={S} k={k} dil={dil}")
            if dil == 1:
                run(lambda: ops.conv1d_tc(xb, W, bias, pad=pad, residual=x, out=y, act2=L, slope2=0.1), flops,
                    f"bf16 c2+res C={C} S={S} k={k}")
    del x, t, y, xs, xb
    torch.cuda.empty_cache()
torch.cuda.synchronize()
