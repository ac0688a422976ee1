"""GPU: 1000 Adam steps (fit.py:28-37,83-113) against the trajectory the unmodified reference produced
(tests/golden/*_long.npz), on the FP32 FMA kernel and -- for the 128-wide nets -- on the tensor-core kernel.

THE BAND.  Two regimes, told apart by how the reference's own arithmetic behaves
(tests/golden/adam_band_oracle.json: the CPU oracle in float32 and in float64 against the same goldens):

* self-adaptive Allen-Cahn cases (ac_sa_small_long, ac_sa_128_long) are well conditioned: the float64 oracle stays
  within 6e-5 of every one of the reference's 1000 losses.  Band: every step's loss within rel 5e-3, the weight
  vector within rel 1e-4 after 100 steps and 1e-3 after 1000 (measured on B200: <= 1.3e-3, 3e-6, 1.8e-5).
* the cases without self-adaptive weights (burgers_long, ac_dist_128_long) amplify rounding: Adam with
  beta_1 = 0.99 rides through loss spikes whose timing shifts with the last bits, and the reference's own numbers
  are not reproducible at that level -- the float32 / float64 oracle departs from the golden by up to 1.06 / 0.77
  (Burgers) and 0.09 / 0.53 (AC) in single steps, 0.09 / 0.064 and 5e-4 / 0.023 in the final weights.  Band: first
  100 steps within rel 0.15 per step and 2e-3 in the weights; over all 1000 steps the MEDIAN per-step deviation
  <= 0.1, the final loss within rel 0.5, and the final weights within rel 0.1 (i.e. a tenth of the distance the
  weights travel, ~0.6-1.0) -- measured: median <= 0.04, final loss <= 0.04, weights <= 0.07.
"""
import os

import numpy as np
import pytest

from tests.adam_band import band_stats
from tests.gpu_util import solver_from_golden

pytestmark = pytest.mark.gpu

WELL = {"loss_rel_max_all": 5e-3, "w_rel_100": 1e-4, "w_rel_end": 1e-3}
SENSITIVE = {"loss_rel_max_100": 0.15, "w_rel_100": 2e-3, "loss_rel_median_all": 0.1, "loss_rel_final": 0.5, "w_rel_end": 0.1}
BAND = {"ac_sa_small_long": WELL, "ac_sa_128_long": WELL, "burgers_long": SENSITIVE, "ac_dist_128_long": SENSITIVE}


@pytest.mark.parametrize("name,kernel", [("ac_sa_small_long", "fma"), ("burgers_long", "fma"),
                                         ("ac_sa_128_long", "fma"), ("ac_sa_128_long", "tc"),
                                         ("ac_dist_128_long", "fma"), ("ac_dist_128_long", "tc")])
def test_adam_1000_steps(name, kernel, monkeypatch):
    if kernel == "fma":
        monkeypatch.setenv("TDQ_DBG", "24")          # FP32 FMA kernel even where the tensor-core kernel applies
    else:
        monkeypatch.delenv("TDQ_DBG", raising=False)
    model, z = solver_from_golden(name)
    model.verbose = False
    n = len(z["adam_loss_hist"])
    assert n == 1000
    model.fit(tf_iter=100)
    w100 = model.u_model.get_flat().copy()
    model.fit(tf_iter=n - 100)
    running = model._engine().term_kernel(model._terms[-1]["tid"])
    assert running.startswith("fused_jet_tc_kernel" if kernel == "tc" else "fused_jet_kernel"), running
    st = band_stats(model.loss_history[:n], w100, model.u_model.get_flat(), z)
    for key, bound in BAND[name].items():
        assert st[key] <= bound, (name, kernel, key, st)
