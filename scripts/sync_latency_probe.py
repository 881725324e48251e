"""Builds (here, `build`) and runs (GPU box) scripts/probes/sync_latency_probe.cu: latencies of mbarrier / tcgen05.commit /
tcgen05.st hand-overs."""
import ctypes, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "probes", "libsync_latency_probe.so")
if "build" in sys.argv or not os.path.exists(SO):
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
                           "-I", os.path.join(HERE, "..", "include"), os.path.join(HERE, "probes", "sync_latency_probe.cu"), "-o", SO,
                           "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    if "build" in sys.argv:
        sys.exit(0)
print("rc", ctypes.CDLL(SO).sync_latency_probe())
