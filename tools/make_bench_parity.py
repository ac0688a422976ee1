#!/usr/bin/env python
"""Writes tests/golden/bench_parity_ac2m.json: what the ORACLE (oracle/pinn_oracle.py, nested autograd on the
CPU, residual mean accumulated over chunks) gets for the first Adam steps of bench.py's workload at its full
size -- Allen-Cahn [2,128x4,1], N_f = 2 000 000 LHS points + IC(512) + periodic(201), Glorot weights seed 0.
bench.py replays the same steps on the GPU(s) after its timed region, at every N, and prints the relative
difference as `parity`: that is how the multi-GPU path is held to the single-device reference numbers on the
driver's own hardware (fit.py:116-190 is the reference path replaced).

    python tools/make_bench_parity.py [--points 2000000] [--steps 3] [--dtype f64|f32]

Minutes of CPU time; the JSON (a few hundred bytes) is committed.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SAMPLE_IDX = [0, 1, 127, 128, 255, 256, 384, 4000, 16895, 16896, 20000, 33408, 40000, 49919, 49920, 50048]


def main():
    import torch
    import bench
    from oracle import pinn_oracle as po
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=bench.N_F)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--chunk", type=int, default=100000)
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    X, ic_X, ic_val, per, w0 = bench.synthetic_problem(args.points)
    up = per[0::2].T.reshape(1, 2, -1, 1).astype(np.float64)
    lo = per[1::2].T.reshape(1, 2, -1, 1).astype(np.float64)
    out = {"workload": "allen_cahn_dist", "n_f": args.points, "steps": args.steps, "sample_idx": SAMPLE_IDX,
           "how": "oracle/pinn_oracle.py loss_and_grads_chunked (chunks of %d points) + KerasAdam, tools/make_bench_parity.py" % args.chunk}
    for tag, dt in (("f64", torch.float64), ("f32", torch.float32)):
        terms = [po.Term("ic", X=ic_X.astype(np.float64), val=ic_val), po.Term("periodic", upper=up, lower=lo, var=["x"])]
        prob = po.Problem(bench.LAYERS, "allen_cahn", X, terms, dtype=dt)
        opt = po.KerasAdam()
        w = w0.copy()
        losses = []
        t0 = time.time()
        for _ in range(args.steps):
            r = po.loss_and_grads_chunked(prob, w, args.chunk)
            losses.append(float(r["loss"]))
            opt.begin_step()
            w = opt.apply("w", w, r["grad"].astype(np.float32))
        out[tag] = {"loss_hist": losses, "w_norm": float(np.linalg.norm(w.astype(np.float64))),
                    "w_sample": [float(w[i]) for i in SAMPLE_IDX], "seconds": time.time() - t0}
        print(tag, losses, out[tag]["w_norm"], "%.0f s" % out[tag]["seconds"], flush=True)
    name = "bench_parity_ac2m.json" if args.points == bench.N_F else "bench_parity_ac_%d.json" % args.points
    with open(os.path.join(ROOT, "tests", "golden", name), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
