"""Developer check: the tensor-core kernel is bit-reproducible across repeated evaluations and fresh engines
(a race in the mbarrier protocol would show up as run-to-run differences)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.test_gpu_properties import ac_engine, points
for n in (100003, 4097, 33):
    X = points(n)
    lam = np.random.RandomState(3).uniform(0, 1, n).astype(np.float32)
    ref = None
    for rep in range(3):
        eng, t = ac_engine(X, lam=lam)
        for it in range(15):
            L, G, _ = eng.loss_grad()
            f = eng.get_residual(t)
            if ref is None:
                ref = (L, G.copy(), f.copy())
            assert L == ref[0] and np.array_equal(G, ref[1]) and np.array_equal(f, ref[2]), (n, rep, it)
        print(n, rep, eng.term_kernel(t), "ok", L)
        eng.close()
print("stress ok")
