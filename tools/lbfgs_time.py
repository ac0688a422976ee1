"""Eager L-BFGS on the self-adaptive Allen-Cahn problem (P = 50 049), 200 iterations timed three times in one process
(tools/gpu_configs.py times the first call only, set-up included)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.gpu_configs import ac
m = ac(50000, True)
for i in range(3):
    t0 = time.perf_counter(); m.fit(tf_iter=0, newton_iter=200); m._engine().sync()
    print("run %d: %.3f ms per iteration" % (i, (time.perf_counter() - t0) * 5))
