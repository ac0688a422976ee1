"""Jet-kernel time against tiles per CTA (Allen-Cahn [2,128x4,1], residual term only + its BC terms): fits
t = a + b * tiles to expose the launch's fixed cost (prologue, pipeline fill / drain, slab epilogue)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from tensordiffeq_b200 import _lib

w_cfg = _lib.Engine.adam_cfg(0.005, 0.99, 0.999, 1e-7, 1)
rows = []
for k in (1, 2, 4, 8, 16, 32, 53, 106):
    n = 148 * 32 * k - 29 * 32          # leaves room for the 29 BC tiles: exactly k tiles on every CTA
    eng, t_res, Xl = bench.build_engine(n, 0, 1, 0)
    for _ in range(5):
        eng.adam_steps(1, w=w_cfg, want_hist=False)
    eng.sync()
    eng.set_kernel_timing(True); eng.kernel_time(t_res, reset=True)
    ms = []
    for _ in range(20):
        eng.mark_begin(); eng.adam_steps(1, w=w_cfg, want_hist=False); ms.append(eng.mark_end())
    k_ms, k_n = eng.kernel_time(t_res)
    rows.append((k, k_ms / k_n * 1e3, float(np.median(ms)) * 1e3))
    print("tiles/CTA %4d   jet kernel %9.1f us   step %9.1f us" % rows[-1], flush=True)
    eng.close()
A = np.array([[1, r[0]] for r in rows]); y = np.array([r[1] for r in rows])
a, b = np.linalg.lstsq(A, y, rcond=None)[0]
print("fit: kernel = %.1f us + %.2f us x tiles" % (a, b))
