"""FP32 pipe probes on the box (developer tool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tensordiffeq_b200 import _lib
eng = _lib.Engine([2, 20, 20, 1])
import sys as _s
for mode in ([int(a) for a in _s.argv[1:]] or (0, 1, 3, 10, 11)):
    print("ffma probe mode", mode, "value %.6g" % eng.ffma_probe(mode), flush=True)
eng.close()
