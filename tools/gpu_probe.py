"""FP32 pipe probes on the box (developer tool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tensordiffeq_b200 import _lib
eng = _lib.Engine([2, 20, 20, 1])
for mode in (0, 1, 2, 3, 4, 1, 3):
    print("ffma probe mode", mode, "TFLOP/s %.2f" % eng.ffma_probe(mode), flush=True)
eng.close()
