"""Debug: per-hand-over clock64() timeline of the panel conv kernel (CTA 0).  Rebuilds with -DSS_PANEL_TRACE on the GPU
box.  Usage: python scripts/panel_trace.py C k [dil]   (SS_PANEL_DBG=15: synchronisation skeleton only)"""
import ctypes, math, os, sys
os.environ["SS_NVCC_EXTRA"] = "-DSS_PANEL_TRACE"
sys.path.insert(0, ".")
from speakingstyle_b200 import build
build.build(force=True)
import numpy as np, torch
from speakingstyle_b200 import ops
C, k = int(sys.argv[1]), int(sys.argv[2])
dil = int(sys.argv[3]) if len(sys.argv) > 3 else 1
B, S = 32, 163840 * 32 // C // 4
x = torch.randn(B, S, C, device="cuda:0")
W = ops.pack_conv(torch.randn(C, C, k, device="cuda:0") / math.sqrt(C * k))
bias = torch.randn(C, device="cuda:0")
t = torch.empty_like(x)
for _ in range(3):
    ops.conv1d(x, W, bias, k=k, dilation=dil, pad=(k - 1) * dil // 2, in_act=ops.ACT_LRELU, in_slope=0.1, out=t, backend="tf32x3_panel")
torch.cuda.synchronize()
buf = np.zeros(8 * 256, dtype=np.int64)
dll = ctypes.CDLL(build.LIB_PATH)
assert dll.ss_debug_panel_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
tr = buf.reshape(8, 256)
t0 = tr[tr > 0].min()
names = ["m:wait<", "m:wait>", "m:commit", "s:ready", "s:slot>", "s:arrive", "s:pan<", "s:pan>"]
print("hand-over " + " ".join(f"{n:>9s}" for n in names))
for i in list(range(0, 30)) + list(range(100, 124)):
    print(f"{i:9d} " + " ".join(f"{int(tr[e, i] - t0):9d}" if tr[e, i] else "        -" for e in range(8)))
d = np.diff(tr[2, 20:200]); d = d[d > 0]
print("MMA thread period per hand-over: median %.0f mean %.0f" % (np.median(d), d.mean()))
