#!/usr/bin/env python
"""bench.py -- collocation points / second per Adam train step (fwd + 2nd-derivative jets + bwd +
optimizer), BASELINE.json's metric, on the Allen-Cahn [2,128x4,1] N_f = 2M configuration
(examples/AC-dist-new.py with N_f raised to 2M as BASELINE.json asks).

  python bench.py --gpus N --steps K --warmup W            # this engine (N>1: under torchrun)
  python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)

A step = one pass of the hot path over the whole synthetic point set: loss + parameter gradients
of every term (residual, IC, periodic), slab reduction, (NCCL all-reduce,) Keras-Adam update.
Strong scaling: N_f is fixed, points are sharded across ranks.  One JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LAYERS = [2, 128, 128, 128, 128, 1]
N_F = 2_000_000
F_ALG = 1_184_256            # algorithmic flop / point: 3 x (fwd matmul flops), C = 4 jets
METRIC = "collocation points/sec per train step (fwd+2nd-deriv+bwd)"
UNIT = "points/s"
L2_FLUSH_BYTES = 256 << 20


def synthetic_problem(n_f, seed=1234):
    """centred LHS points on x in [-1,1], t in [0,1]; IC on 512 points; periodic couples on a
    bootstrap resample of 201 time levels; Glorot weights, seed 0."""
    from tensordiffeq_b200.synthetic import centred_lhs, glorot_normal_flat
    X = centred_lhs(n_f, [[-1, 1], [0, 1]], seed).astype(np.float32)
    x_ic = np.linspace(-1, 1, 512)
    ic_X = np.stack([x_ic, np.zeros_like(x_ic)], 1).astype(np.float32)
    ic_val = (x_ic ** 2 * np.cos(math.pi * x_ic)).astype(np.float32).reshape(-1, 1)
    t_lv = np.linspace(0, 1, 201)[np.random.RandomState(7).randint(0, 201, 201)]
    per = np.empty((402, 2), np.float32)
    per[0::2, 0], per[1::2, 0] = 1.0, -1.0
    per[0::2, 1] = per[1::2, 1] = t_lv
    w0 = glorot_normal_flat(LAYERS, seed=0)
    return X, ic_X, ic_val, per, w0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-i", str(gpu_index), "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def count(self):
        """samples written so far"""
        try:
            self.f.flush()
            return sum(1 for _ in open(self.f.name))
        except Exception:
            return 0

    def stop(self, skip=0):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines()[skip:]:
            c = [v.strip() for v in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def build_engine(n_f, rank, world, local_rank, stream=None, sa=False):
    from tensordiffeq_b200 import _lib, tf
    from tensordiffeq_b200 import trace as T
    from tensordiffeq_b200.models import _trace, _integrand

    def f_model(u_model, x, t):                 # examples/AC-dist-new.py:38-46
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)
        u_xx = tf.gradients(u_x, x)
        u_t = tf.gradients(u, t)
        c1 = tf.constant(.0001)
        c2 = tf.constant(5.0)
        return u_t - c1 * u_xx + c2 * u * u * u - c2 * u

    X, ic_X, ic_val, per, w0 = synthetic_problem(n_f)
    eng = _lib.Engine(LAYERS, device=local_rank, stream=stream)
    fe = T.as_sym(_trace(f_model, 2))
    # sa: per-point self-adaptive weights on the residual and the IC (examples/AC-SA.py:49-56), Adam ascent on them
    ops, slots, orders = T.lower({"out": _integrand(fe, False, "mse" if sa else None), "f": fe}, 2)
    t_res = eng.add_term(2, orders, ops, slots["out"], slots["f"], 0)
    u = T.jet((0, 0))
    ops, slots, orders = T.lower({"out": _integrand(u, True, "mse" if sa else None)}, 2)
    t_ic = eng.add_term(2, orders, ops, slots["out"], -1, 0)
    ops, slots, orders = T.lower({"out": u}, 2)
    t_per = eng.add_term(2, orders, ops, slots["out"], -1, 1)
    base, rem = divmod(n_f, world)
    s = rank * base + min(rank, rem)
    c = base + (1 if rank < rem else 0)
    Xl = np.ascontiguousarray(X[s:s + c])
    eng.set_points(t_res, Xl, n_f)
    # BC / IC terms are counted once; each lives whole on one rank, spread round-robin so that no
    # single rank arrives late at the all-reduce
    if rank == t_ic % world:
        eng.set_points(t_ic, ic_X, len(ic_X))
        eng.set_target(t_ic, ic_val)
        if sa:
            eng.set_lambda(t_ic, (100 * np.random.RandomState(2).uniform(size=len(ic_X))).astype(np.float32))
    else:
        eng.set_points(t_ic, ic_X[:0], len(ic_X))
    if rank == t_per % world:
        eng.set_points(t_per, per, len(per) // 2)
    else:
        eng.set_points(t_per, per[:0], len(per) // 2)
    if sa:
        eng.set_lambda(t_res, np.random.RandomState(1).uniform(size=n_f).astype(np.float32)[s:s + c])
    eng.set_params(w0)
    return eng, t_res, Xl


PARITY_STEPS = 3


def parity_check(eng, n_f, w_cfg):
    """After the timed region: restart from the seed-0 weights, take the first Adam steps again and compare
    the loss of every step and the weights after the last one with what the CPU oracle got for the same
    workload at its full size (tests/golden/bench_parity_ac2m.json, written by tools/make_bench_parity.py,
    float64 accumulation).  Runs at every N: with sharded points and the all-reduce in the path this is the
    check that the multi-GPU step computes the single-device numbers (fit.py:116-190)."""
    from tensordiffeq_b200.synthetic import glorot_normal_flat
    name = "bench_parity_ac2m.json" if n_f == N_F else "bench_parity_ac_%d.json" % n_f
    eng.set_params(glorot_normal_flat(LAYERS, seed=0))
    eng.adam_reset()
    hist = [float(v) for v in eng.adam_steps(PARITY_STEPS, w=w_cfg, want_hist=True)]
    w = eng.get_params().astype(np.float64)
    out = {"steps": PARITY_STEPS, "loss_hist": hist, "w_norm": float(np.linalg.norm(w)), "against": None, "rel": None}
    try:
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", name)))
    except Exception:
        out["note"] = "no committed oracle values for N_f=%d (tools/make_bench_parity.py --points %d writes them)" % (n_f, n_f)
        return out
    g = gold["f64"]
    idx = gold["sample_idx"]
    ws, gs = w[idx], np.asarray(g["w_sample"], np.float64)
    rel_loss = max(abs(a - b) / abs(b) for a, b in zip(hist, g["loss_hist"][:PARITY_STEPS]))
    rel_norm = abs(out["w_norm"] - g["w_norm"]) / g["w_norm"]
    rel_samp = float(np.linalg.norm(ws - gs) / np.linalg.norm(gs))
    out.update(against="tests/golden/%s (CPU oracle, float64 accumulation, %d Adam steps from the seed-0 weights)" % (name, PARITY_STEPS),
               rel_loss_max=rel_loss, rel_w_norm=rel_norm, rel_w_sample=rel_samp, w_sample=[float(v) for v in ws],
               rel=max(rel_loss, rel_norm, rel_samp), tolerance=1e-4)
    return out


def measure_sa(n_f, rank, world, local_rank, steps, flush, dist):
    """The same step with per-point self-adaptive weights (BASELINE config 2's arithmetic: lambda, its two Adam
    moments and its gradient stream through the residual kernel, fused lambda-Adam ascent in its epilogue):
    32 algorithmic HBM bytes per point instead of 8."""
    import torch
    from tensordiffeq_b200 import _lib
    eng, t_res, Xl = build_engine(n_f, rank, world, local_rank, sa=True)
    if world > 1:
        uid = [eng.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, 0)
        eng.comm_init(uid[0], rank, world)
    w_cfg = _lib.Engine.adam_cfg(0.005, 0.99, 0.999, 1e-7, 1)
    l_cfg = _lib.Engine.adam_cfg(0.005, 0.99, 0.999, 1e-7, -1)
    for _ in range(3):
        eng.adam_steps(1, w=w_cfg, lam=l_cfg, want_hist=False)
    eng.sync()
    eng.set_kernel_timing(True)
    eng.kernel_time(t_res, reset=True)
    ms = []
    for _ in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        eng.mark_begin()
        eng.adam_steps(1, w=w_cfg, lam=l_cfg, want_hist=False)
        ms.append(eng.mark_end())
    k_ms, k_n = eng.kernel_time(t_res, reset=True)
    tot = float(np.sum(ms))
    if world > 1:
        tt = torch.tensor([tot], dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        tot = float(tt.item())
    kern_ms = k_ms / max(k_n, 1)
    n_local = len(Xl)
    eng.close()
    return {"n_f": n_f, "steps": steps, "ms_per_step": tot / steps, "value": n_f / (tot / steps * 1e-3), "kernel_ms": kern_ms,
            "hbm_bytes_per_point": 32, "hbm_achieved_gbs": n_local * 32 / (kern_ms * 1e-3) / 1e9,
            "note": "points 8 B + lambda read/write 8 B + its Adam moments read/write 16 B per point; compute-bound like the plain step"}


def run_ours(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this engine has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=world)
    from tensordiffeq_b200 import _lib
    n_f = args.points
    eng, t_res, Xl = build_engine(n_f, rank, world, local_rank)
    if world > 1:
        uid = [eng.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, 0)
        eng.comm_init(uid[0], rank, world)
    w_cfg = _lib.Engine.adam_cfg(0.005, 0.99, 0.999, 1e-7, 1)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        eng.sync()
        if world > 1:
            dist.barrier()

    for _ in range(max(args.warmup, 3)):
        eng.adam_steps(1, w=w_cfg, want_hist=False)
    barrier()
    ffma_peak = 0.0
    if rank == 0:
        try:                                  # measured FFMA rate: developer probe library (tools/probe), optional
            from tools import probe as _probe
            ffma_peak = _probe.ffma_peak(local_rank)
        except Exception:
            ffma_peak = 0.0
    barrier()

    # ---- timed region: K steps, device time per step (CUDA events on the engine's stream), L2 flushed
    #      between steps (not timed), max over ranks
    eng.set_kernel_timing(True)
    eng.kernel_time(t_res, reset=True)
    launches0 = eng.info()["launches"]
    sampler = ClockSampler(local_rank) if rank == 0 else None
    s0 = 0
    if sampler:
        time.sleep(0.35)                 # let nvidia-smi start polling
        s0 = sampler.count()
    if world > 1:
        dist.barrier()
    wall0 = time.perf_counter()
    step_ms = []
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        eng.mark_begin()
        eng.adam_steps(1, w=w_cfg, want_hist=False)
        step_ms.append(eng.mark_end())
    barrier()
    wall = time.perf_counter() - wall0
    launches_timed = eng.info()["launches"] - launches0
    # nvidia-smi polls every 100 ms: when the timed region is shorter than that (multi-GPU strong scaling),
    # keep running the same step, untimed, until the sampler has seen the GPU under this load
    k_ms, k_n = eng.kernel_time(t_res, reset=True)
    eng.set_kernel_timing(False)
    extra = 0
    t_x = time.perf_counter()
    while True:
        cont = 1 if (sampler and sampler.count() - s0 < 2 and time.perf_counter() - t_x < 5.0) else 0
        if world > 1:                                    # the step holds a collective: every rank follows rank 0
            tt = torch.tensor([cont], dtype=torch.int32)
            dist.broadcast(tt, 0)
            cont = int(tt.item())
        if not cont:
            break
        eng.adam_steps(1, w=w_cfg, want_hist=False)
        eng.sync()
        extra += 1
    clocks = sampler.stop(skip=s0) if sampler else None
    if clocks is not None:
        clocks["window"] = "timed region" if extra == 0 else "timed region + %d identical untimed steps right after it" % extra
    launches = launches_timed
    total_ms = float(np.sum(step_ms))
    if world > 1:
        tt = torch.tensor([total_ms], dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms = float(tt.item())
    ms_per_step = total_ms / args.steps
    value = n_f / (ms_per_step * 1e-3)

    # ---- end to end through the C-ABI with HOST buffers: every step uploads the rank's point set from
    #      pinned host memory, runs the step, reads the loss back
    pinned = torch.from_numpy(Xl).pin_memory()
    Xp = pinned.numpy()
    e2e_steps = max(3, min(args.steps, 10))
    barrier()
    e2e_ms = []
    for i in range(e2e_steps + 1):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.set_points(t_res, Xp, n_f)
        h = eng.adam_steps(1, w=w_cfg, want_hist=True)
        _ = float(h[0])
        eng.sync()
        if i:
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
    e2e_tot = float(np.sum(e2e_ms))
    if world > 1:
        tt = torch.tensor([e2e_tot], dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_tot = float(tt.item())
    e2e_value = n_f / (e2e_tot / e2e_steps * 1e-3)
    # ... and the same with the upload of the NEXT step's point set running on the engine's copy stream while this step
    # computes (tdq_prefetch_term_points / tdq_swap_term_points: points resampled between steps): every timed step
    # still moves its 16 MB from pinned host memory and reads its loss back
    barrier()
    st_ms = []
    for i in range(e2e_steps + 1):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.prefetch_points(t_res, Xp)
        h = eng.adam_steps(1, w=w_cfg, want_hist=True)
        _ = float(h[0])
        eng.swap_points(t_res)
        eng.sync()
        if i:
            st_ms.append((time.perf_counter() - t0) * 1e3)
    st_tot = float(np.sum(st_ms))
    if world > 1:
        tt = torch.tensor([st_tot], dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        st_tot = float(tt.item())
    e2e_streamed = n_f / (st_tot / e2e_steps * 1e-3)

    parity = parity_check(eng, n_f, w_cfg)
    sa = None
    if not args.no_sa:
        # the lambda streams (north_star: "achieved HBM GB/s on the point and lambda streams")
        sa = {"n_f_%d" % n_f: measure_sa(n_f, rank, world, local_rank, max(3, min(args.steps, 10)), flush, dist)}
        if world == 1 and n_f != 50000:
            sa["n_f_50000"] = measure_sa(50000, rank, world, local_rank, 20, flush, dist)   # examples/AC-SA.py:12
    if rank != 0:
        return
    n_local = len(Xl)
    kern_ms = k_ms / max(k_n, 1)
    achieved = n_local * F_ALG / (kern_ms * 1e-3) / 1e12 if k_n else None
    info = eng.info()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sm_max = float(peaks.get("sm_max_mhz", 1965.0))
    nominal = 2 * 128 * info["n_sm"] * sm_max * 1e6 / 1e12
    kernel_name = eng.term_kernel(t_res)
    on_tensor = kernel_name.startswith("fused_jet_tc_kernel")
    # Tensor ceiling in fp32-equivalent TFLOP/s.  Every fp32 product costs one kind::tf32 MMA (hi.hi; dense tf32 runs at half
    # the measured bf16 rate = 2 bf16-units) plus two bf16 correction MMAs (hi.lo + lo.hi, 1 unit each): 4 bf16-units per
    # product, so ceiling = bf16 rate / 4.  (Round 1 and the 16-point kernel run all three terms as tf32: rate / 6.)
    # The kernel is timed launch by launch with CUDA events, with an L2 flush (GPU otherwise idle) between launches and
    # the SM clock at its maximum: the burst figure is the denominator; the sustained one is reported beside it.
    bf16_burst = float(peaks.get("bf16_tflops", 2250.0 * 0.73))
    bf16_sust = float(peaks.get("bf16_tflops_sustained", bf16_burst))
    units = 6.0 if "tc16" in kernel_name else 4.0
    tensor_peak = bf16_burst / units
    tensor_peak_sust = bf16_sust / units
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    alg_bytes = n_local * 2 * 4                          # point stream: 4 B x d per point (no lambdas here)
    # DRAM bytes per launch and tensor-pipe activity come from the committed `ncu --set full` capture of this kernel
    # at this size (profiles/r02_traffic.json); at other sizes / rank counts the keys are left out rather than guessed
    traffic = pipe_active = smem_pipe = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        if tj.get("n_f") == n_local:
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
            pipe_active = tj.get("tensor_pipe_active_pct")
            if tj.get("smem_pipe_tc_pct") is not None and tj.get("smem_pipe_lsu_pct") is not None:
                smem_pipe = {"tensor_core_operand_reads_pct": tj["smem_pipe_tc_pct"], "worker_loads_stores_pct": tj["smem_pipe_lsu_pct"],
                             "note": "share of the shared-memory data pipe's peak over the whole launch (same ncu capture): the resource "
                                     "that bounds this kernel -- each phase of a tile is ~65 % bound by it, DESIGN.md sections 5 and 9"}
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 (hidden-layer GEMMs as a split fp32 product on tcgen05: tf32 x tf32 plus two bf16 correction terms, fp32 accumulate; everything else fp32)" if on_tensor else "f32", "data": "synthetic",
        "config": {"workload": "allen_cahn_dist [2,128x4,1] tanh, N_f=%d LHS points + IC(512) + periodic(201), Adam step" % n_f,
                   "n_f": n_f, "points_per_gpu": n_local, "l2": "256 MiB buffer written between timed steps (flush)",
                   "timing": "CUDA events per step on the engine stream, sum over K steps, max over ranks"},
        "roofline": {"bound": "tensor" if on_tensor else "fp32", "achieved": achieved,
                     "peak": tensor_peak if on_tensor else nominal, "unit": "TFLOP/s",
                     "frac": (achieved / (tensor_peak if on_tensor else nominal)) if achieved else None, "traffic": traffic,
                     "traffic_note": "dram__bytes_read+write of one launch (profiles/r02_traffic.json, ncu --set full); algorithmic HBM bytes are %d" % alg_bytes,
                     "tensor_pipe_active_pct": pipe_active, "smem_pipe": smem_pipe,
                     "peak_sustained": tensor_peak_sust if on_tensor else None,
                     "frac_of_peak_sustained": (achieved / tensor_peak_sust) if (achieved and on_tensor) else None,
                     "peak_source": ("split fp32 product on tcgen05 (1 tf32 MMA + 2 bf16 correction MMAs = %d bf16-units): MEASURED_PEAKS.json "
                                     "bf16_tflops (burst) %.1f / %d; achieved counts fp32-equivalent algorithmic flops; "
                                     "against round 1's all-tf32 ceiling (rate / 6) the same achieved figure reads %.3f"
                                     % (int(units), bf16_burst, int(units), (achieved / (bf16_burst / 6.0)) if achieved else 0.0)) if on_tensor else
                                    "nominal FFMA: 2 x 128 lanes x %d SMs x %.0f MHz (MEASURED_PEAKS.json has no fp32 entry)" % (info["n_sm"], sm_max),
                     "fp32_fma_peak": nominal, "frac_of_fp32_fma_peak": (achieved / nominal) if achieved else None,
                     "ffma_measured_tflops": ffma_peak,
                     "frac_of_ffma_measured": (achieved / ffma_peak) if (achieved and ffma_peak) else None,
                     "kernel": kernel_name, "kernel_ms": kern_ms, "kernel_launches_timed": k_n,
                     "kernel_share_of_step": kern_ms / ms_per_step if ms_per_step else None,
                     "flop_per_point": F_ALG,
                     "hbm_stream": {"achieved_gbs": alg_bytes / (kern_ms * 1e-3) / 1e9 if k_n else None, "peak_gbs": hbm_peak,
                                    "note": "point stream only (8 B/pt): the path is compute-bound by design"}},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(Xp.nbytes), "d2h_bytes_per_step": 4,
                "steps": e2e_steps, "note": "tdq_set_term_points (pinned host -> device) + tdq_adam_steps(1) + loss read-back per step, strictly one after the other",
                "streamed": {"value": e2e_streamed, "unit": UNIT, "h2d_bytes_per_step": int(Xp.nbytes), "d2h_bytes_per_step": 4, "steps": e2e_steps,
                             "note": "tdq_prefetch_term_points (the next step's points, pinned host -> device on the copy stream) + tdq_adam_steps(1) + loss read-back + tdq_swap_term_points per step: the upload overlaps the step"}},
        "gpu_launches": int(launches),
        "comm": eng.comm_mode(),
        "parity": parity,
        "ac_sa": sa,
        "clocks": clocks,
        "wall_s_timed_region": wall,
    }
    if world > 1:
        line["roofline"].pop("traffic", None)        # measured on one GPU at the single-GPU size only
        line["roofline"].pop("traffic_note", None)
        line["roofline"].pop("tensor_pipe_active_pct", None)
        line["roofline"].pop("smem_pipe", None)
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(sample_points=args.cpu_points, steps=3, warmup=1)
    elif not args.no_cpu_baseline:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)


def cpu_step_fn(n_pts, threads=None):
    """the oracle's restatement of one reference Adam step (nested autograd + Keras Adam) on n_pts points"""
    import torch
    from oracle import pinn_oracle as po
    if threads:
        torch.set_num_threads(threads)
    X, ic_X, ic_val, per, w0 = synthetic_problem(n_pts)
    up = per[0::2].T.reshape(1, 2, -1, 1).astype(np.float64)
    lo = per[1::2].T.reshape(1, 2, -1, 1).astype(np.float64)
    terms = [po.Term("ic", X=ic_X.astype(np.float64), val=ic_val),
             po.Term("periodic", upper=up, lower=lo, var=["x"])]
    prob = po.Problem(LAYERS, "allen_cahn", X, terms)
    opt = po.KerasAdam()
    state = {"w": w0.copy()}

    def step():
        r = po.loss_and_grads(prob, state["w"])
        opt.begin_step()
        state["w"] = opt.apply("w", state["w"], r["grad"])
        return float(r["loss"])
    return step


def cpu_baseline(sample_points, steps, warmup):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step = cpu_step_fn(sample_points, cores)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return {"value": sample_points / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d of the %d collocation points per step (same net, same terms), %d timed steps; "
                      "torch-CPU restatement of the reference step (TensorFlow is not installable here)" % (sample_points, N_F, steps),
            "s_per_step": dt}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    n = args.cpu_points
    step = cpu_step_fn(n, cores)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    v = n / dt
    sample = ("%d of the %d collocation points per step; oracle port of the reference step "
              "(nested autograd + Keras Adam) on %d host threads" % (n, N_F, cores))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "allen_cahn_dist [2,128x4,1] tanh, N_f=%d (bounded sample of %d points per step)" % (N_F, n)},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)      # >= 100 timed steps
    ap.add_argument("--warmup", type=int, default=10)     # ... after >= 10 warm-up steps
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--points", type=int, default=N_F)
    ap.add_argument("--cpu-points", type=int, default=50000)
    ap.add_argument("--no-sa", action="store_true", help="skip the self-adaptive (lambda stream) measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = min(args.steps, 20)
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
