"""Micro-benchmark of single convolution launches (run on the GPU box).  usage: tc_microbench.py [case ...]"""
import math
import sys
import time

import torch

sys.path.insert(0, ".")
from speakingstyle_b200 import ops  # noqa: E402

DEV = "cuda:0"
CASES = {
    # name: (B, S, Cin, Cout, k, dil, kind)
    "hifi128_c1": (32, 58816, 128, 128, 11, 3, "c1"),
    "hifi128_c2": (32, 58816, 128, 128, 11, 1, "c2"),
    "hifi32_c1": (32, 235264, 32, 32, 3, 1, "c1"),
    "ffn_w1": (64, 696, 256, 1024, 9, 1, "plain"),
    "qkv": (64, 696, 256, 768, 1, 1, "plain"),
}


def bench(name, reps=5):
    B, S, Cin, Cout, k, dil, kind = CASES[name]
    x = torch.randn(B, S, Cin, device=DEV).to(torch.bfloat16)
    W = ops.pack_conv(torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k))
    bias = torch.randn(Cout, device=DEV)
    res = torch.randn(B, S, Cout, device=DEV) if kind == "c2" else None
    pad = (k - 1) // 2 * dil

    def go():
        if kind == "c1":
            ops.conv1d_tc(x, W, bias, dilation=dil, pad=pad, want_y=False, act2=ops.ACT_LRELU, slope2=0.1)
        elif kind == "c2":
            ops.conv1d_tc(x, W, bias, dilation=dil, pad=pad, residual=res, act2=ops.ACT_LRELU, slope2=0.1)
        else:
            ops.conv1d_tc(x, W, bias, dilation=dil, pad=pad)

    for _ in range(2):
        go()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        go()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    fl = 2.0 * B * S * Cout * k * Cin
    print(f"{name:12s} {ms:8.3f} ms  {fl / ms / 1e9:8.1f} TFLOP/s  out-elems {B * S * Cout / 1e6:.0f} M")


if __name__ == "__main__":
    for n in (sys.argv[1:] or list(CASES)):
        bench(n, reps=1 if len(sys.argv) > 1 else 5)
