"""LengthRegulator kernel alone against measured HBM bandwidth (run on the GPU box)."""
import json
import sys

import torch

sys.path.insert(0, ".")
import bench  # noqa: E402

torch.cuda.set_device(0)
print(json.dumps(bench.lr_roofline(torch.device("cuda", 0), bench.peaks())))
