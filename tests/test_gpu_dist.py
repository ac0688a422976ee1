"""GPU, 2 ranks (skipped on a single-GPU box): dist=True through the public API -- points sharded across
two processes / two GPUs, BC terms on rank 0, one packed NCCL all-reduce inside the library -- must
reproduce the reference's SINGLE-device loss, gradient and Adam trajectory."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, q, no_peer):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), TDQ_NO_PEER=str(no_peer))
    from tests.gpu_util import solver_from_golden, rel
    model, z = solver_from_golden(name, dist=True)
    model.verbose = False
    loss, grads = model.grad()
    n_net = 2 * (len(z["layer_sizes"]) - 1)
    flat = np.concatenate([g.ravel() for g in grads[:n_net]])
    out = {"rank": rank, "loss": float(loss), "grad_rel": rel(flat, z["grad0"]), "loss_ref": float(z["loss0"]),
           "comm": model._engine().comm_mode()}
    if "lamgrad0_0" in z:
        out["lam_rel"] = rel(grads[n_net], z["lamgrad0_0"])
    n = len(z["adam_loss_hist"])
    model.fit(tf_iter=n)
    out["hist"] = np.array(model.loss_history[:n], np.float32)
    out["hist_ref"] = z["adam_loss_hist"]
    out["w_rel"] = rel(model.u_model.get_flat(), z["w_after_adam"])
    if "lam_after_adam_0" in z:
        out["lam_after_rel"] = rel(model.lambdas[0].numpy(), z["lam_after_adam_0"])
    q.put(out)


@pytest.mark.parametrize("no_peer", [0, 1], ids=["nvlink-peer", "nccl"])
@pytest.mark.parametrize("name", ["ac_dist_128", "ac_sa_small"])
def test_two_rank_training_matches_single_device_reference(name, no_peer):
    """both exchange paths: the fused reduce + NVLink peer exchange + Adam kernel (when the box allows peer mapping), and
    reduce -> ncclAllReduce -> Adam"""
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, q, no_peer)) for r in range(2)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=300) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for o in outs:
        assert abs(o["loss"] - o["loss_ref"]) <= 1e-4 * abs(o["loss_ref"]), o
        assert o["grad_rel"] < 1e-4, o
        assert np.allclose(o["hist"], o["hist_ref"], rtol=1e-3), o
        assert o["w_rel"] < 1e-3, o
        if "lam_rel" in o:
            assert o["lam_rel"] < 1e-4 and o["lam_after_rel"] < 1e-4, o
    assert outs[0]["loss"] == outs[1]["loss"]          # replicas stay bit-identical
    assert outs[0]["comm"] == outs[1]["comm"]
    if no_peer:
        assert outs[0]["comm"].startswith("nccl")
