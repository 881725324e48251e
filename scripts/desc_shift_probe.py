"""Builds and runs scripts/probes/desc_shift_probe.cu: does a swizzled K-major tcgen05 operand accept a start address
at an arbitrary row of the panel, and with which base_offset?  `python scripts/desc_shift_probe.py build` here (nvcc),
`python scripts/desc_shift_probe.py` on the GPU box."""
import ctypes
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "probes", "libdesc_shift_probe.so")
if "build" in sys.argv or not os.path.exists(SO):
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-shared",
                           "-Xcompiler", "-fPIC", "-I", os.path.join(HERE, "..", "include"),
                           os.path.join(HERE, "probes", "desc_shift_probe.cu"), "-o", SO, "-lcudart_static", "-ldl", "-lrt",
                           "-lpthread"])
    if "build" in sys.argv:
        sys.exit(0)
import torch  # noqa: E402

NS = 40
out = torch.zeros(2, NS, 128, 32, device="cuda:0")
lib = ctypes.CDLL(SO)
rc = lib.desc_shift_probe(ctypes.c_void_p(out.data_ptr()))
print("rc", rc)
o = out.cpu()
i = torch.arange(128).view(1, 128, 1)
n = torch.arange(32).view(1, 1, 32)
for v in range(2):
    ok = []
    for s in range(NS):
        want = ((s + i) * 8 + n // 4).float()[0]
        ok.append(bool(torch.equal(o[v, s], want)))
    print("variant", v, "(base_offset = s & 7)" if v else "(base_offset = 0)", "exact for shifts:",
          [s for s in range(NS) if ok[s]])
    bad = [s for s in range(NS) if not ok[s]]
    if bad:
        s = bad[0]
        print("  first bad shift", s, "row0", o[v, s, 0, ::4].tolist(), "row1", o[v, s, 1, ::4].tolist())
