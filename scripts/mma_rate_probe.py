"""Builds (here, `build`) and runs (GPU box) scripts/probes/mma_rate_probe.cu: cycles per small-N tf32 tcgen05.mma."""
import ctypes, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "probes", "libmma_rate_probe.so")
if "build" in sys.argv or not os.path.exists(SO):
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
                           "-I", os.path.join(HERE, "..", "include"), os.path.join(HERE, "probes", "mma_rate_probe.cu"), "-o", SO,
                           "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    if "build" in sys.argv:
        sys.exit(0)
print("rc", ctypes.CDLL(SO).mma_rate_probe())
