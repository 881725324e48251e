"""Builds (here, `build`) and runs (GPU box) scripts/probes/handover_probe.cu: cycles per operand hand-over of the TS-form
pipeline skeleton.  Run under `timeout`."""
import ctypes, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "probes", "libhandover_probe.so")
if "build" in sys.argv or not os.path.exists(SO):
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
                           "-I", os.path.join(HERE, "..", "include"), os.path.join(HERE, "probes", "handover_probe.cu"), "-o", SO,
                           "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    if "build" in sys.argv:
        sys.exit(0)
print("rc", ctypes.CDLL(SO).handover_probe())
