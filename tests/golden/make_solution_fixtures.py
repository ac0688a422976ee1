#!/usr/bin/env python
"""Reference-held solution fields -> small .npz fixtures (run in the build container only):
  /root/reference/examples/AC.mat            -> tests/golden/ac_solution.npz       (x[512], t[201], u[512,201])
  /root/reference/examples/burgers_shock.mat -> tests/golden/burgers_solution.npz  (x[256], t[100], u[256,100])
The example scripts compare `model.predict` with these fields (examples/AC-SA.py:68-89,
examples/burgers-new.py:44-68); tools/accuracy.py does the same on the GPU box, where /root/reference
does not exist."""
import os

import numpy as np
import scipy.io

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/examples"

d = scipy.io.loadmat(os.path.join(REF, "AC.mat"))
np.savez_compressed(os.path.join(HERE, "ac_solution.npz"), x=d["x"].ravel().astype(np.float64),
                    t=d["tt"].ravel().astype(np.float64), u=np.real(d["uu"]).astype(np.float32))
d = scipy.io.loadmat(os.path.join(REF, "burgers_shock.mat"))
np.savez_compressed(os.path.join(HERE, "burgers_solution.npz"), x=d["x"].ravel().astype(np.float64),
                    t=d["t"].ravel().astype(np.float64), u=np.real(d["usol"]).astype(np.float32))
for n in ("ac_solution", "burgers_solution"):
    z = np.load(os.path.join(HERE, n + ".npz"))
    print(n, {k: z[k].shape for k in z.files})
