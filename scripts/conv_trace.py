"""Debug: per-K-chunk clock64() timeline of the 3xTF32 conv kernels (CTA 0).  Rebuilds with -DSS_CONV_TRACE on
the GPU box.  Usage: python scripts/conv_trace.py [ss|ts]"""
import ctypes, math, os, sys
os.environ["SS_NVCC_EXTRA"] = "-DSS_CONV_TRACE -DSS_CONV2_TRACE"
sys.path.insert(0, ".")
from speakingstyle_b200 import build
build.build(force=True)
import numpy as np, torch
from speakingstyle_b200 import ops
var = sys.argv[1] if len(sys.argv) > 1 else "ts"
B, S, Cin, Cout, k = int(os.environ.get("TRACE_B", "64")), int(os.environ.get("TRACE_S", "696")), 256, 1024, 9
if len(sys.argv) > 2:
    Cin, Cout, k = (int(v) for v in sys.argv[2:5])
x = torch.randn(B, S, Cin, device="cuda:0")
W = ops.pack_conv(torch.randn(Cout, Cin, k, device="cuda:0") / math.sqrt(Cin * k))
bias = torch.randn(Cout, device="cuda:0")
cold = os.environ.get("TRACE_COLD", "0") == "1"
res = torch.randn(B, S, Cout, device="cuda:0") if os.environ.get("TRACE_RES", "0") == "1" else None
for _ in range(3):
    ops.conv1d(x, W, bias, pad=k // 2, act=ops.ACT_RELU, backend="tf32x3_" + var, residual=res)
if cold:  # evict x / residual / weights from L2 before the traced launch
    junk = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0"); junk.zero_(); junk.add_(1)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); ops.conv1d(x, W, bias, pad=k // 2, act=ops.ACT_RELU, backend="tf32x3_" + var, residual=res); b.record(); torch.cuda.synchronize()
print(f"{var}: {a.elapsed_time(b):.3f} ms (trace build)")
NEV = 12 if var in ("ts", "ts2") else 8
buf = np.zeros(NEV * 256, dtype=np.int64)
dll = ctypes.CDLL(build.LIB_PATH)
fn = getattr(dll, f"ss_debug_conv_{var}_trace")
assert fn(buf.ctypes.data_as(ctypes.c_void_p)) == 0
t = buf.reshape(NEV, 256)
t0 = t[0, 0]
names = ["tma", "full", "xf<", "mma>", "mma<", "st?", "st>", "epi<", "m:cemp", "m:temp", "e:cfull", "e:cemp"][:NEV]
print("chunk " + " ".join(f"{n:>7s}" for n in names))
NCH = int(os.environ.get("TRACE_ROWS", "124"))
for i in list(range(0, 12)) + list(range(NCH - 24, NCH)):
    print(f"{i:5d} " + " ".join(f"{int(t[e, i] - t0):7d}" if t[e, i] else "      -" for e in range(NEV)))
hi = min(240, int((t[4] != 0).sum()))
d = np.diff(t[4, 8:hi])
print("MMA issue period per chunk (clk): mean %.0f median %.0f" % (d.mean(), np.median(d)))
print("tma->full latency: median %.0f" % np.median((t[1] - t[0])[8:hi]))
print("full->xf<: median %.0f" % np.median((t[2] - t[1])[8:hi]))
