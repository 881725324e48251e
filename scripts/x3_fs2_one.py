"""The decoder FFT block's four convolutions on the split-bf16 kernel at the cfg2 shape (B = 64, S = 696, d_model 256,
d_inner 1024; transformer/SubLayers.py:39-55, 85-93), timed or - with --ncu - launched once each for
`ncu -k regex:bf16x3`."""
import math, sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
B, S, D, DI = 64, 696, 256, 1024
x = torch.randn(B, S, D, device="cuda:0")
xp = ops.split_bf16(x)
mk = lambda co, ci, k: ops.pack_conv(torch.randn(co, ci, k, device="cuda:0") / math.sqrt(ci * k))
Wqkv, Wfc, W1, W2 = mk(3 * D, D, 1), mk(D, D, 1), mk(DI, D, 9), mk(D, DI, 1)
bq, bf, b1, b2 = (torch.randn(n, device="cuda:0") for n in (3 * D, D, DI, D))
hp = torch.empty(B, S, 2, DI, dtype=torch.bfloat16, device="cuda:0")
cases = [
    ("QKV 256->768 k=1", lambda: ops.conv1d_x3(xp, Wqkv, bq), 2.0 * B * S * D * 3 * D),
    ("fc 256->256 k=1 + residual", lambda: ops.conv1d_x3(xp, Wfc, bf, residual=x), 2.0 * B * S * D * D),
    ("w_1 256->1024 k=9 -> relu pair", lambda: ops.conv1d_x3(xp, W1, b1, pad=4, want_y=False, act2=ops.ACT_RELU, out_pair=hp),
     2.0 * B * S * D * DI * 9),
    ("w_2 1024->256 k=1 + residual", lambda: ops.conv1d_x3(hp, W2, b2, residual=x), 2.0 * B * S * D * DI),
]
for name, fn, flops in cases:
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    if "--ncu" in sys.argv:
        continue
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(f"{name:36s} {ms * 1e3:8.1f} us  {flops / ms / 1e9:7.1f} TFLOP/s", flush=True)
