"""Debug: per-tile clock64() timeline of the tcgen05 attention kernels' pipelines (CTA 0).
Rebuilds the library with -DSS_ATT_TRACE on the GPU box (the shipped build has no tracing).
Usage: python scripts/attn_trace.py [32|128]"""
import ctypes, os, sys
os.environ["SS_NVCC_EXTRA"] = "-DSS_ATT_TRACE"
sys.path.insert(0, ".")
from speakingstyle_b200 import build
build.build(force=True)
import numpy as np, torch
from speakingstyle_b200 import ops
dk = int(sys.argv[1]) if len(sys.argv) > 1 else 128
B, S, H, BKV = (64, 696, 2, 32) if dk == 128 else (64, 800, 8, 64)
qkv = torch.randn(B, S, 3 * H * dk, device="cuda:0")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda:0")
for _ in range(3):
    ops.attention(qkv, lens, H, backend="tc")
torch.cuda.synchronize()
buf = np.zeros(16 * 64, dtype=np.int64)
dll = ctypes.CDLL(build.LIB_PATH)
fn = dll.ss_debug_att_trace if dk == 128 else dll.ss_debug_att32_trace
rc = fn(buf.ctypes.data_as(ctypes.c_void_p))
assert rc == 0, rc
t = buf.reshape(16, 64)
NT = (S + BKV - 1) // BKV
t0 = t[0, 0]
names = ["tmaV", "tmaK", "xfK>", "xfK<", "xfV>", "xfV<", "S>", "S<", "PV>", "PV<", "smx>", "P<", "fold>", "fold<", "exp<"]
print(f"dk={dk}  tile " + " ".join(f"{n:>7s}" for n in names))
for j in range(NT):
    print(f"{j:4d} " + " ".join(f"{int(t[e, j] - t0):7d}" if t[e, j] else "      -" for e in range(15)))
d = np.diff(t[9, :NT])
print("PV issue period per tile (clk): mean %.0f  median %.0f" % (d.mean(), np.median(d)))
