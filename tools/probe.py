"""ctypes binding of tools/probe/libtdq_probe.so: the developer micro-benchmarks (FFMA / FFMA2 rates, tcgen05
3xTF32 rate and accuracy, operand-layout experiments) behind the design decisions in DESIGN.md.  Not part of the
product library.  python tools/probe.py [mode ...]   (mode list: tools/probe/probe.cu, tdq_probe)"""
import ctypes as C
import os
import sys

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "probe", "libtdq_probe.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH)
        _lib.tdq_probe.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_float)]
        _lib.tdq_probe_ffma_peak.argtypes = [C.c_int, C.POINTER(C.c_float)]
    return _lib


def probe(mode, device=0):
    v = C.c_float(0)
    rc = load().tdq_probe(device, int(mode), C.byref(v))
    if rc:
        raise RuntimeError("tdq_probe(%d) failed (%d)" % (mode, rc))
    return v.value


def ffma_peak(device=0):
    """measured FP32 FFMA throughput of the device in TFLOP/s"""
    v = C.c_float(0)
    if load().tdq_probe_ffma_peak(device, C.byref(v)):
        raise RuntimeError("tdq_probe_ffma_peak failed")
    return v.value


if __name__ == "__main__":
    for mode in ([int(a) for a in sys.argv[1:]] or (0, 1, 3, 10, 11)):
        print("probe mode", mode, "value %.6g" % probe(mode), flush=True)
