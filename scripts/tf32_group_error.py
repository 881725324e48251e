"""RMS / max error of the 3xTF32 conv against float64 as a function of the accumulator-restart group."""
import math, sys, torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
torch.manual_seed(0)
for (Cin, Cout, k) in ((256, 1024, 9), (1024, 256, 1), (1024, 1024, 3)):
    B, S = 2, 300
    x = torch.randn(B, S, Cin, device=DEV)
    w = torch.randn(Cout, Cin, k, device=DEV) / math.sqrt(Cin * k)
    W = ops.pack_conv(w)
    ref = F.conv1d(x.double().transpose(1, 2), w.double(), padding=k // 2).transpose(1, 2)
    scale = ref.pow(2).mean().sqrt().item()
    def err(y):
        d = (y.double() - ref)
        return d.pow(2).mean().sqrt().item() / scale, d.abs().max().item() / scale
    print(f"K={Cin * k}: ffma rms %.2e max %.2e" % err(ops.conv1d(x, W, None, pad=k // 2, backend="ffma")))
    for var in ("ss", "ts"):
        for g in (2, 4, 8, 16, 1000):
            ops.tf32_k_group = g
            print(f"   {var} group={g:4d}: rms %.2e max %.2e" % err(ops.conv1d(x, W, None, pad=k // 2, backend="tf32x3_" + var)))
ops.tf32_k_group = 0
