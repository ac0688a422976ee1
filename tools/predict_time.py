"""Device time of tdq_predict's kernel on 2M points (AC [2,128x4,1]): run under ncu --metrics gpu__time_duration.sum"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from tensordiffeq_b200 import _lib, tf
from tensordiffeq_b200 import trace as T
from tensordiffeq_b200.models import _trace
from tests.test_jet_host import f_ac
X, ic_X, ic_val, per, w0 = bench.synthetic_problem(2_000_000)
eng = _lib.Engine(bench.LAYERS)
eng.set_params(w0)
fe = T.as_sym(_trace(f_ac, 2))
ops, slots, orders = T.lower({"f": fe}, 2)
for _ in range(3):
    u, f = eng.predict(2, orders, ops, slots["f"], slots["f"], X)
print("predict ok", float(np.abs(f).mean()))
