"""TEST INFRASTRUCTURE ONLY: CPU oracle for the text -> mel -> wave path.

fs2_oracle.py   functional torch-CPU restatement of the reference arithmetic
intops.c        plain-C restatement of the integer-valued steps (loaded through `intops()`)
ref_harness.py  imports the real reference from /root/reference (build container only)
make_golden.py  writes tests/golden/*.npz from the real reference

Nothing in speakingstyle_b200/ imports this package.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libss_oracle.so")
_handle = None


def build_intops(force=False):
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(
            os.path.join(_HERE, "intops.c")):
        subprocess.check_call(["make", "-s", "-C", _HERE] + (["-B"] if force else []))
    return _LIB


def intops():
    """ctypes handle on the C oracle (built on demand)."""
    global _handle
    if _handle is None:
        lib = ctypes.CDLL(build_intops())
        i64, f32p, i64p = ctypes.c_int64, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int64)
        lib.ss_oracle_round_durations.argtypes = [f32p, ctypes.c_float, i64, f32p]
        lib.ss_oracle_bucketize.argtypes = [f32p, i64, f32p, ctypes.c_int, i64p]
        lib.ss_oracle_length_regulate.argtypes = [f32p, f32p, i64, i64, i64, i64, f32p, i64p, i64p]
        lib.ss_oracle_length_regulate.restype = i64
        lib.ss_oracle_mel_len.argtypes = [f32p, i64, i64, ctypes.POINTER(ctypes.c_int32)]
        lib.ss_oracle_wave_int16.argtypes = [f32p, i64, ctypes.c_float, ctypes.POINTER(ctypes.c_int16)]
        _handle = lib
    return _handle
