"""CPU, world_size 2: the host logic of the multi-GPU path; gloo stands in for NCCL.  Each rank takes
its contiguous shard of the collocation points (models._Dist.shard), evaluates its partial loss /
gradient scaled by 1/N_f(global) -- with the ORACLE standing in for the device kernels, there is no
GPU here -- rank 0 alone adds the BC/IC terms, and a SUM all-reduce of the packed
[grad | loss] vector must reproduce the single-process result.  The same decomposition is
exercised on one real GPU by tests/test_gpu_properties.py::test_shard_additivity_and_tail_200k, and
the NCCL path by bench.py --gpus N."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import pinn_oracle as po
from tests import golden_util as gu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, out_q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    from tensordiffeq_b200.models import _Dist
    d = _Dist(True)
    payload = d.exchange_bytes(b"unique-id-from-rank-0" if rank == 0 else b"")   # the NCCL id path (tensordiffeq_b200/rendezvous.py)
    assert payload == b"unique-id-from-rank-0"
    got = d.gather_shards(np.full(3 + rank, float(rank), np.float32))            # the lambda read-back path
    assert np.array_equal(got, np.concatenate([np.zeros(3, np.float32), np.ones(4, np.float32)]))
    dist.init_process_group("gloo", rank=rank, world_size=world)                 # stands in for the NCCL all-reduce below
    z = gu.load(name)
    n = len(z["X_f"])
    s, c = d.shard(n)
    terms = gu.terms_of(z) if rank == 0 else []           # BC/IC terms are counted once
    prob = po.Problem([int(v) for v in z["layer_sizes"]], gu.RESIDUAL_OF[name], z["X_f"][s:s + c], terms)
    w = torch.tensor(z["w0"], requires_grad=True)
    loss, _ = po.update_loss(prob, w)
    # update_loss averages the residual over the LOCAL points: rescale that part to the global mean
    prob_bc = po.Problem(prob.layer_sizes, prob.f_model, z["X_f"][s:s + 1], terms)
    res_only = po.Problem(prob.layer_sizes, prob.f_model, z["X_f"][s:s + c], [])
    l_res, _ = po.update_loss(res_only, w)
    l_bc = loss - l_res
    local = l_res * (c / n) + l_bc
    (g,) = torch.autograd.grad(local, w)
    vec = torch.cat([g.detach(), local.detach().reshape(1)])
    dist.all_reduce(vec, op=dist.ReduceOp.SUM)            # the packed message of tdq_comm_init's all-reduce
    if rank == 0:
        out_q.put(vec.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["ac_dist_128"])
def test_sharded_sum_equals_single_process(name):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, q)) for r in range(2)]
    for p in procs:
        p.start()
    vec = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    z = gu.load(name)
    assert abs(vec[-1] - z["loss0"]) <= 1e-5 * abs(z["loss0"])
    assert np.linalg.norm(vec[:-1] - z["grad0"]) / np.linalg.norm(z["grad0"]) < 1e-5
