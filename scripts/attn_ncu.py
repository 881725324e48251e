"""One launch of each tcgen05 attention kernel at the cfg2 shapes (for ncu captures)."""
import sys, torch
sys.path.insert(0, ".")
from speakingstyle_b200 import ops
DEV = "cuda:0"
for (B, S, H, dk) in ((64, 800, 8, 32), (64, 696, 2, 128)):
    D = H * dk
    qkv = torch.randn(B, S, 3 * D, device=DEV)
    lens = torch.full((B,), S, dtype=torch.int32, device=DEV)
    for _ in range(2):
        ops.attention(qkv, lens, H, backend="tc")
    torch.cuda.synchronize()
