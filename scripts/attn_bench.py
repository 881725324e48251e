import sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
for (B, S, H, dk) in ((64, 800, 8, 32), (16, 400, 8, 32), (64, 696, 2, 128), (64, 119, 2, 128)):
    D = H * dk
    qkv = torch.randn(B, S, 3 * D, device=DEV)
    lens = torch.full((B,), S, dtype=torch.int32, device=DEV)
    for be in ("simt", "tc"):
        for _ in range(2): ops.attention(qkv, lens, H, backend=be)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5): ops.attention(qkv, lens, H, backend=be)
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        print(f"B={B} S={S} H={H} dk={dk} {be:5s} {ms:8.3f} ms  {4.0 * S * S * D * B / ms / 1e9:7.1f} TFLOP/s")
