"""Statistics of a long Adam trajectory against the reference's (tests/golden/*_long.npz, 1000 steps of the
unmodified reference package): the quantities DESIGN.md section 2 states a band for."""
import numpy as np


def band_stats(loss_hist, w100, w_end, z):
    g = np.asarray(z["adam_loss_hist"], np.float64)
    h = np.asarray(loss_hist, np.float64)[:len(g)]
    r = np.abs(h - g) / np.abs(g)

    def rel(a, b):
        a = np.asarray(a, np.float64).ravel()
        b = np.asarray(b, np.float64).ravel()
        return float(np.linalg.norm(a - b) / np.linalg.norm(b))
    return {"loss_rel_max_100": float(r[:100].max()), "loss_rel_max_all": float(r.max()),
            "loss_rel_median_all": float(np.median(r)), "loss_rel_final": float(r[-1]),
            "w_rel_100": rel(w100, z["w_after_100"]), "w_rel_end": rel(w_end, z["w_after_adam"]),
            "w_move_100": rel(z["w_after_100"], z["w0"]), "w_move_end": rel(z["w_after_adam"], z["w0"])}
