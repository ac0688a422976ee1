"""CPU: host-side logic of the product -- the C-ABI library loads and exports every symbol the header
declares, the boundary / domain builders reproduce what the reference built (golden fixtures), the
lowering of a whole problem into terms, argument checking with the reference's exception strings.
No compute call is made (no GPU here); the CUDA path is covered by the -m gpu tests."""
import math
import os
import re

import numpy as np
import pytest

import tensordiffeq_b200 as tdq
from tensordiffeq_b200 import _lib, tf
from tensordiffeq_b200 import trace as T
from tensordiffeq_b200.boundaries import DomainND, IC, dirichletBC, periodicBC, FunctionDirichletBC, FunctionNeumannBC
from tensordiffeq_b200.models import CollocationSolverND, DiscoveryModel, _Dist
from tests import golden_util as gu
from tests.gpu_util import solver_from_golden, deriv_model
from tests.test_jet_host import f_ac, f_burgers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "tdq_b200.h")).read()
    declared = set(re.findall(r"\b(tdq_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"tdq_handle"}
    lib = _lib.load()
    bound = {name for name, _, _ in _lib.SYMBOLS}
    assert declared == bound, (declared ^ bound)
    for name in declared:
        assert getattr(lib, name) is not None


def test_bindings_agree_with_the_header_prototypes():
    """every ctypes signature in _lib.SYMBOLS has the header's parameter count, and the same width wherever the header
    says int32_t / int64_t / float (pointers and handles are only checked to be pointers)"""
    import ctypes as C
    hdr = open(os.path.join(ROOT, "include", "tdq_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    protos = {m.group(2): (m.group(1).strip(), m.group(3)) for m in re.finditer(r"\n\s*([A-Za-z_ ]+?\**)\s*(tdq_[a-z_0-9]+)\s*\(([^;{]*?)\)\s*;", hdr)}
    scalars = {"int32_t": C.c_int32, "int64_t": C.c_int64, "float": C.c_float, "int": C.c_int}
    for name, res, args in _lib.SYMBOLS:
        assert name in protos, name
        params = [a.strip() for a in protos[name][1].split(",") if a.strip() and a.strip() != "void"]
        assert len(params) == len(args), (name, params, args)
        for prm, ct in zip(params, args):
            if "*" in prm:
                assert ct in (C.c_void_p, C.c_char_p) or hasattr(ct, "contents") or issubclass(ct, C._Pointer), (name, prm, ct)
            else:
                base = prm.replace("const", "").split()[0]
                assert scalars[base] is ct, (name, prm, ct)


def test_no_cpu_path():
    # in this container there is no GPU: creating an engine must fail loudly, not fall back
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.TdqError, match="no CUDA device"):
        _lib.Engine([2, 20, 20, 1])


def test_boundaries_match_reference_construction():
    """same seed, same call order as tests/golden/make_golden.py:ac_problem -> same point sets"""
    z = gu.load("ac_sa_small")
    np.random.seed(11)
    D = DomainND(["x", "t"], time_var="t")
    D.add("x", [-1.0, 1.0], 100)
    D.add("t", [0.0, 1.0], 50)
    D.generate_collocation_points(2000)
    init = IC(D, [lambda x: x ** 2 * np.cos(math.pi * x)], var=[["x"]])
    per = periodicBC(D, ["x"], [deriv_model])
    assert np.allclose(init.input, z["bc0_input"])
    assert np.allclose(np.asarray(init.val), z["bc0_val"], atol=1e-7)
    assert np.allclose(np.asarray(per.upper), z["bc1_upper"])
    assert np.allclose(np.asarray(per.lower), z["bc1_lower"])
    # LHS: centred cells, every cell hit exactly once per column
    X = D.X_f
    assert X.shape == (2000, 2)
    cells = np.sort(np.floor((X[:, 0] + 1) / 2 * 2000).astype(int))
    assert np.array_equal(cells, np.arange(2000))


def test_poisson_boundaries_match_reference_construction():
    z = gu.load("poisson")
    np.random.seed(5)
    D = DomainND(["x", "y"])
    D.add("x", [0, 1.0], 11)
    D.add("y", [0, 1.0], 11)
    D.generate_collocation_points(100)
    fx = lambda y: -tf.math.sin(tf.constant(math.pi) * y) * tf.math.sin(tf.constant(math.pi))  # noqa: E731
    lower_x = dirichletBC(D, val=0.0, var="x", target="upper")
    upper_x = FunctionDirichletBC(D, fun=[fx], var="x", target="upper", func_inputs=["y"], n_values=10)
    upper_y = FunctionDirichletBC(D, fun=[fx], var="y", target="upper", func_inputs=["x"], n_values=10)
    lower_y = dirichletBC(D, val=0.0, var="y", target="lower")
    for i, bc in enumerate([upper_x, lower_x, upper_y, lower_y]):
        assert np.allclose(bc.input, z["bc%d_input" % i]), i
        assert np.allclose(np.asarray(bc.val, np.float32), z["bc%d_val" % i], atol=1e-7), i


def test_term_lowering_ac_sa():
    model, z = solver_from_golden("ac_sa_small")
    descs = model._build_term_descs()
    assert [d["kind"] for d in descs] == ["ic", "periodic", "residual"]
    ic, per, res = descs
    assert ic["lam"] == 1 and res["lam"] == 0            # lambda order: residual first, then BCs (utils.py:99-112)
    assert per["pair"] == 1 and per["X"].shape == (2 * 50, 2)
    assert np.all(per["X"][0::2, 0] == 1.0) and np.all(per["X"][1::2, 0] == -1.0)
    assert np.array_equal(per["X"][0::2, 1], per["X"][1::2, 1])
    ops, slots, orders = T.lower({"out": per["out"]}, 2)
    assert orders == [0, 0]                               # quirk Q2: only u of (u, u_x) enters the periodic loss
    ops, slots, orders = T.lower({"out": res["out"], "f": res["f"]}, 2)
    assert orders == [2, 1] and any(o[0] == "LAMBDA" for o in ops)


def test_term_lowering_residual_tuple():
    """a callback returning two residuals becomes ONE term over the shared points (models.py:160-188)"""
    for name, n_lam_sq in (("ac_two_res", 1), ("ac_two_res_both", 2)):
        model, z = solver_from_golden(name)
        descs = model._build_term_descs()
        assert [d["kind"] for d in descs] == ["ic", "periodic", "residual"]
        res = descs[-1]
        assert res["lam"] == 0 and len(model._f_exprs) == 2
        ops, slots, orders = T.lower({"out": res["out"], "f": res["f"]}, 2)
        assert orders == [2, 2]                           # union of u_xx (first residual) and u_tt (second)
        # numeric check of the merged integrand against the two residuals evaluated apart
        rng = np.random.RandomState(0)
        jets = rng.standard_normal((5, 7))
        coords = rng.standard_normal((2, 7))
        lam = rng.uniform(size=7)
        out = T.eval_program(ops, jets, coords, lam=lam)[slots["out"]]
        u, u_x, u_xx, u_t, u_tt = jets[0], jets[1], jets[2], jets[3], jets[4]
        f1 = u_t - np.float32(.0001) * u_xx + 5.0 * u ** 3 - 5.0 * u
        f2 = u_tt - u * u_x + 0.25 * coords[0]
        want = (lam * f1) ** 2 + ((lam * f2) ** 2 if n_lam_sq == 2 else f2 ** 2)
        assert np.allclose(out, want, rtol=1e-6)


def test_mixed_partial_lowering():
    """tf.gradients(tf.gradients(u, x), y) (examples/testing.py:23-31) becomes the jet channel after the pure ones; higher
    mixed derivatives are refused"""
    from tests.gpu_util import f_mixed
    from tensordiffeq_b200.models import _trace
    fe = T.as_sym(_trace(f_mixed, 2))
    ops, slots, orders = T.lower({"f": fe}, 2)
    assert list(orders) == [2, 2] and orders.mixed == 1
    assert any(o[0] == "JET" and o[1] == 5 for o in ops)          # u, u_x, u_xx, u_y, u_yy, u_xy
    rs = np.random.RandomState(0)
    jets, coords = rs.standard_normal((6, 5)), rs.uniform(0, 1, (2, 5))
    u, u_x, u_xx, u_y, u_yy, u_xy = jets
    want = u_xx + 0.5 * u_xy * u + u_yy + u_xy * u_xy - np.sin(np.float32(math.pi) * coords[0]) * np.sin(np.float32(math.pi) * coords[1])
    assert np.allclose(T.eval_program(ops, jets, coords)[slots["f"]], want, rtol=1e-6)

    def f_xxy(u_model, x, y):
        u = u_model(tf.concat([x, y], 1))
        return tf.gradients(tf.gradients(tf.gradients(u, x)[0], x)[0], y)[0]
    with pytest.raises(T.MixedPartialError):
        _trace(f_xxy, 2)


def test_reference_exception_strings():
    D = DomainND(["x", "t"], time_var="t")
    D.add("x", [-1.0, 1.0], 16)
    D.add("t", [0.0, 1.0], 8)
    D.generate_collocation_points(64)
    init = IC(D, [lambda x: x], var=[["x"]])
    per = periodicBC(D, ["x"], [deriv_model])
    m = CollocationSolverND(verbose=False)
    m.compile([2, 8, 1], f_ac, D, [init, per], isAdaptive=True,
              dict_adaptive={"residual": [False], "BCs": [False, True]},
              init_weights={"residual": [None], "BCs": [None, tf.random.uniform([8, 1])]})
    with pytest.raises(Exception, match="not accepting Adapative Periodic Boundaries Conditions"):
        m._build_term_descs()
    m.compile([2, 8, 1], f_ac, D, [init, per], dist=True)
    with pytest.raises(Exception, match="Currently we dont support distributed minibatching training"):
        m.fit(tf_iter=1, batch_sz=16)
    with pytest.raises(Exception, match="Assimilate needs to be set to 'true'"):
        m.compile_data(None, None, None)
    with pytest.raises(AttributeError):
        # quirk Q10: initialize_weights_loss(None) throws before the "no inputs were specified" check
        CollocationSolverND().compile([2, 8, 1], f_ac, D, [init], isAdaptive=True)


def test_shard_partition():
    import os as _os
    tot = 0
    for r in range(3):
        _os.environ.update(RANK=str(r), WORLD_SIZE="3", LOCAL_RANK=str(r))
        s, c = _Dist(True).shard(10)
        assert s == tot
        tot += c
    assert tot == 10
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        _os.environ.pop(k)


def test_network_layout_and_save_load(tmp_path):
    net = tdq.networks.neural_net([2, 5, 3, 1], seed=3)
    flat = net.get_flat()
    ws = net.get_weights()
    assert [w.shape for w in ws] == [(2, 5), (5,), (5, 3), (3,), (3, 1), (1,)]
    assert np.array_equal(np.concatenate([w.ravel() for w in ws]), flat)     # utils.py:19-29 order
    assert all(np.all(b == 0) for b in ws[1::2])                              # keras bias init
    net.save(str(tmp_path / "m"))
    net2 = tdq.networks.NeuralNet.load(str(tmp_path / "m"))
    assert np.array_equal(net2.get_flat(), flat)
    sw, sb = tdq.utils.get_sizes([2, 5, 3, 1])
    assert sw == [10, 15, 3] and sb == [5, 3, 1]


def test_discovery_compile_traces_vars():
    params = [tf.Variable(1e-4), tf.Variable(1e-4)]
    from tests.test_jet_host import f_ch
    from tensordiffeq_b200.models import _trace
    for i, v in enumerate(params):
        v._sym = T.var(i)
    fe = T.as_sym(_trace(f_ch, 2, extra_first=params))
    for v in params:
        v._sym = None
    ops, slots, orders = T.lower({"f": fe}, 2)
    assert orders == [4, 1]
    assert sorted({o[1] for o in ops if o[0] == "VAR"}) == [0, 1]
    assert len(ops) <= _lib.MAX_OPS


def test_rendezvous_three_ranks_over_tcp():
    """tensordiffeq_b200/rendezvous.py (what dist=True uses between ranks instead of a framework): broadcast from rank 0
    and all-gather in rank order, three ranks as threads on the loopback interface"""
    import socket
    import threading
    from tensordiffeq_b200.rendezvous import Rendezvous
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = {}

    def run(rank):
        r = Rendezvous(rank, 3, addr="127.0.0.1", port=port, timeout=30.0)
        got = r.broadcast(b"nccl-unique-id" if rank == 0 else b"")
        parts = r.all_gather(bytes([rank]) * (rank + 1))
        r.barrier()
        r.close()
        out[rank] = (got, parts)
    th = [threading.Thread(target=run, args=(k,)) for k in range(3)]
    for t in th:
        t.start()
    for t in th:
        t.join(60)
    assert all(not t.is_alive() for t in th)
    for k in range(3):
        assert out[k][0] == b"nccl-unique-id"
        assert out[k][1] == [b"\x00", b"\x01\x01", b"\x02\x02\x02"]
