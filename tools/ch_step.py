"""BASELINE config 4 (Cahn-Hilliard discovery, [2,128x4,1], 25 600 points) as a loop of Adam steps: the workload the
16-point-tile tensor-core kernel is profiled on (ncu -k regex:fused_jet_tc16 ... python tools/ch_step.py)."""
import math, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tensordiffeq_b200 import tf
from tensordiffeq_b200.models import DiscoveryModel
from tests.test_jet_host import f_ch

x = np.linspace(-1, 1, 256); t = np.linspace(0, 1, 100); X, T = np.meshgrid(x, t)
Xs = np.hstack((X.flatten()[:, None], T.flatten()[:, None]))
u = np.cos(math.pi * Xs[:, 0:1]) * np.exp(-Xs[:, 1:2])
params = [tf.Variable(1e-4), tf.Variable(1e-4)]
dm = DiscoveryModel(); dm.compile([2, 128, 128, 128, 128, 1], f_ch, [Xs[:, 0:1], Xs[:, 1:2]], tf.constant(u), params)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dm.train_op(5); dm._engine().sync()
eng = dm._engine()
eng.set_kernel_timing(True); eng.kernel_time(dm._t_res, reset=True)
t0 = time.perf_counter(); dm.train_op(n); eng.sync(); dt = (time.perf_counter() - t0) / n
k_ms, k_n = eng.kernel_time(dm._t_res)
print("CH discovery: %.3f ms/step, jet kernel %.3f ms (%s), %.1f TFLOP/s fp32-equivalent (1 775 616 flop/point)" % (
    dt * 1e3, k_ms / max(k_n, 1), eng.term_kernel(dm._t_res), 25600 * 1775616 / (k_ms / max(k_n, 1) * 1e-3) / 1e12))
