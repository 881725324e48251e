"""Numerics probe for the 3xTF32 tcgen05 kernel (run on the GPU box): error vs fp64 as a function of
the reduction length, for random fp32 operands, TF32-exact operands (isolates the accumulation) and
all-positive operands (exposes a rounding bias)."""
import math
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from speakingstyle_b200 import ops  # noqa: E402

DEV = "cuda:0"


def run(Cin, k, mode, S=512, Cout=128):
    g = torch.Generator().manual_seed(Cin * 31 + k)
    x = torch.randn(1, S, Cin, generator=g)
    w = torch.randn(Cout, Cin, k, generator=g) / math.sqrt(Cin * k)
    if mode == "tf32exact":
        x, w = ops.tf32_round(x), ops.tf32_round(w)
    if mode == "positive":
        x, w = x.abs(), w.abs()
    ref = F.conv1d(x.transpose(1, 2).double(), w.double(), None, padding=(k - 1) // 2).transpose(1, 2)
    W = ops.pack_conv(w.to(DEV))
    out = {}
    for be in ("ffma", "tf32x3"):
        y = ops.conv1d(x.to(DEV), W, None, pad=(k - 1) // 2, backend=be).cpu().double()
        e = y - ref
        out[be] = (e.pow(2).mean().sqrt().item(), e.abs().max().item(), (e / ref.abs().clamp_min(1e-3)).mean().item())
    scale = ref.abs().mean().item()
    print(f"K={Cin * k:5d} ({Cin}x{k}) {mode:10s} |ref|~{scale:.2f}  ffma rms {out['ffma'][0]:.2e} max {out['ffma'][1]:.2e} "
          f"bias {out['ffma'][2]:+.1e} | tf32x3 rms {out['tf32x3'][0]:.2e} max {out['tf32x3'][1]:.2e} bias {out['tf32x3'][2]:+.1e}")


for mode in ("random", "tf32exact", "positive"):
    for Cin, k in ((32, 1), (256, 1), (256, 3), (256, 9), (1024, 3)):
        run(Cin, k, mode)
