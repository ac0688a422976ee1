"""ctypes binding of libtdq_b200.so (C-ABI: include/tdq_b200.h).

There is no CPU fallback: if the shared library is missing or no B200 is visible
the product path raises."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtdq_b200.so")

OP = dict(CONST=0, JET=1, COORD=2, TARGET=3, LAMBDA=4, VAR=5, ADD=6, SUB=7, MUL=8, DIV=9, NEG=10, SIN=11, COS=12,
          EXP=13, TANH=14, SQRT=15, LOG=16, POWI=17, POWF=18)
MAX_OPS = 160
MAX_VARS = 8
MAX_DIMS = 3
MAX_ORDER = 4


class tdq_op(C.Structure):
    _fields_ = [("op", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("imm", C.c_float)]


class tdq_term_spec(C.Structure):
    _fields_ = [("n_dims", C.c_int32), ("orders", C.c_int32 * MAX_DIMS), ("ops", C.POINTER(tdq_op)),
                ("n_ops", C.c_int32), ("out_slot", C.c_int32), ("f_slot", C.c_int32), ("pair", C.c_int32), ("mixed", C.c_int32)]


class tdq_adam_cfg(C.Structure):
    _fields_ = [("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("sign", C.c_int32)]


_FP = C.POINTER(C.c_float)
_H = C.c_void_p

# every symbol include/tdq_b200.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("tdq_create", C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_void_p, C.POINTER(_H)]),
    ("tdq_destroy", C.c_int, [_H]),
    ("tdq_last_error", C.c_char_p, [_H]),
    ("tdq_info", C.c_int, [_H, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ("tdq_set_params", C.c_int, [_H, _FP, C.c_int64]),
    ("tdq_get_params", C.c_int, [_H, _FP, C.c_int64]),
    ("tdq_set_param_mask", C.c_int, [_H, _FP, C.c_int64]),
    ("tdq_set_vars", C.c_int, [_H, _FP, C.c_int32]),
    ("tdq_get_vars", C.c_int, [_H, _FP, C.c_int32]),
    ("tdq_add_term", C.c_int, [_H, C.POINTER(tdq_term_spec), C.POINTER(C.c_int32)]),
    ("tdq_set_term_points", C.c_int, [_H, C.c_int32, _FP, C.c_int64, C.c_int64]),
    ("tdq_prefetch_term_points", C.c_int, [_H, C.c_int32, _FP, C.c_int64]),
    ("tdq_swap_term_points", C.c_int, [_H, C.c_int32]),
    ("tdq_set_term_range", C.c_int, [_H, C.c_int32, C.c_int64, C.c_int64, C.c_int64]),
    ("tdq_set_term_target", C.c_int, [_H, C.c_int32, _FP, C.c_int64]),
    ("tdq_set_term_lambda", C.c_int, [_H, C.c_int32, _FP, C.c_int64]),
    ("tdq_get_term_lambda", C.c_int, [_H, C.c_int32, _FP, C.c_int64]),
    ("tdq_get_term_lambda_grad", C.c_int, [_H, C.c_int32, _FP, C.c_int64]),
    ("tdq_get_term_residual", C.c_int, [_H, C.c_int32, _FP, C.c_int64]),
    ("tdq_loss_grad", C.c_int, [_H, _FP, _FP, _FP]),
    ("tdq_adam_steps", C.c_int, [_H, C.POINTER(tdq_adam_cfg), C.POINTER(tdq_adam_cfg), C.POINTER(tdq_adam_cfg),
                                 C.c_int32, _FP]),
    ("tdq_adam_epochs", C.c_int, [_H, C.POINTER(tdq_adam_cfg), C.POINTER(tdq_adam_cfg), C.POINTER(tdq_adam_cfg),
                                  C.POINTER(C.c_int32), C.c_int32, C.c_int64, C.c_int32, C.c_int32, _FP]),
    ("tdq_adam_reset", C.c_int, [_H]),
    ("tdq_lbfgs", C.c_int, [_H, C.c_int32, C.c_float, _FP, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    ("tdq_lbfgs_hz", C.c_int, [_H, C.c_int32, C.c_float, _FP, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    ("tdq_predict", C.c_int, [_H, C.POINTER(tdq_term_spec), _FP, C.c_int64, _FP, _FP]),
    ("tdq_loss_grad_static", C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.POINTER(tdq_term_spec), C.c_void_p, C.c_void_p,
                                       C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32]),
    ("tdq_comm_unique_id", C.c_int, [C.c_void_p]),
    ("tdq_comm_init", C.c_int, [_H, C.c_void_p, C.c_int32, C.c_int32]),
    ("tdq_comm_mode", C.c_int, [_H, C.c_char_p, C.c_int32]),
    ("tdq_mark_begin", C.c_int, [_H]),
    ("tdq_mark_end", C.c_int, [_H, _FP]),
    ("tdq_sync", C.c_int, [_H]),
    ("tdq_kernel_time", C.c_int, [_H, C.c_int32, C.c_int32, _FP, C.POINTER(C.c_int64)]),
    ("tdq_term_kernel", C.c_int, [_H, C.c_int32, C.c_char_p, C.c_int32]),
    ("tdq_set_kernel_timing", C.c_int, [_H, C.c_int32]),
]

_lib = None


def load():
    """dlopen the library and type every entry point.  Raises if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    global LIB_PATH
    LIB_PATH = os.environ.get("TDQ_LIB", LIB_PATH)        # developer builds (e.g. -DTDQ_TC_PROFILE) live elsewhere
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "tensordiffeq_b200: %s is missing -- build it with `make -C tensordiffeq_b200/csrc` "
            "(or __graft_entry__.build()).  There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def fptr(a):
    return a.ctypes.data_as(_FP)


def f32c(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class TdqError(Exception):
    pass


def make_spec(n_dims, orders, ops, out_slot, f_slot=-1, pair=0):
    """ops: list of (opname|int, a, b, imm).  Returns (spec, keepalive)."""
    arr = (tdq_op * len(ops))()
    for i, (op, a, b, imm) in enumerate(ops):
        arr[i].op = OP[op] if isinstance(op, str) else int(op)
        arr[i].a = int(a)
        arr[i].b = int(b)
        arr[i].imm = float(imm)
    spec = tdq_term_spec()
    spec.n_dims = n_dims
    for i in range(MAX_DIMS):
        spec.orders[i] = int(orders[i]) if i < len(orders) else 0
    spec.ops = C.cast(arr, C.POINTER(tdq_op))
    spec.n_ops = len(ops)
    spec.out_slot = out_slot
    spec.f_slot = f_slot
    spec.pair = pair
    spec.mixed = int(getattr(orders, "mixed", 0))     # trace.Orders: the program also reads u_xy
    return spec, arr


class Engine:
    """Thin object wrapper over one tdq_handle."""

    def __init__(self, layer_sizes, device=0, stream=None):
        self.lib = load()
        self.h = _H()
        ls = (C.c_int32 * len(layer_sizes))(*[int(v) for v in layer_sizes])
        rc = self.lib.tdq_create(ls, len(layer_sizes), int(device), C.c_void_p(stream) if stream else None, C.byref(self.h))
        if rc:
            raise TdqError(self.lib.tdq_last_error(None).decode())
        self.layer_sizes = [int(v) for v in layer_sizes]
        self.n_params = sum(self.layer_sizes[i - 1] * self.layer_sizes[i] + self.layer_sizes[i]
                            for i in range(1, len(self.layer_sizes)))
        self.d = self.layer_sizes[0]
        self._keep = []
        self.term_n = {}

    def _ck(self, rc):
        if rc:
            raise TdqError(self.lib.tdq_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.lib.tdq_destroy(self.h)
            self.h = _H()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters
    def set_params(self, w):
        w = f32c(w).ravel()
        self._ck(self.lib.tdq_set_params(self.h, fptr(w), w.size))

    def get_params(self):
        w = np.empty(self.n_params, np.float32)
        self._ck(self.lib.tdq_get_params(self.h, fptr(w), w.size))
        return w

    def set_param_mask(self, mask):
        m = f32c(mask).ravel()
        self._ck(self.lib.tdq_set_param_mask(self.h, fptr(m), m.size))

    def set_vars(self, v):
        v = f32c(v).ravel()
        self.n_vars = v.size
        self._ck(self.lib.tdq_set_vars(self.h, fptr(v) if v.size else None, v.size))

    def get_vars(self, n):
        v = np.empty(n, np.float32)
        self._ck(self.lib.tdq_get_vars(self.h, fptr(v) if n else None, n))
        return v

    # ---- terms
    def add_term(self, n_dims, orders, ops, out_slot, f_slot=-1, pair=0):
        spec, keep = make_spec(n_dims, orders, ops, out_slot, f_slot, pair)
        tid = C.c_int32(-1)
        self._ck(self.lib.tdq_add_term(self.h, C.byref(spec), C.byref(tid)))
        return tid.value

    def set_points(self, term, X, n_global=None):
        X = f32c(X)
        n = X.shape[0]
        self.term_n[term] = n
        self._ck(self.lib.tdq_set_term_points(self.h, term, fptr(X) if n else None, n, int(n_global or max(n, 1))))

    def prefetch_points(self, term, X):
        """asynchronous upload of the term's NEXT point set (same count) while steps run on the current one"""
        X = f32c(X)
        self._staged = getattr(self, "_staged", {})
        self._staged[term] = X            # the host buffer must outlive the copy
        self._ck(self.lib.tdq_prefetch_term_points(self.h, term, fptr(X), X.shape[0]))

    def swap_points(self, term):
        self._ck(self.lib.tdq_swap_term_points(self.h, term))
        getattr(self, "_staged", {}).pop(term, None)

    def set_range(self, term, start, count, n_global):
        self._ck(self.lib.tdq_set_term_range(self.h, term, int(start), int(count), int(n_global)))

    def set_target(self, term, t):
        t = f32c(t).ravel()
        self._ck(self.lib.tdq_set_term_target(self.h, term, fptr(t), t.size))

    def set_lambda(self, term, lam):
        lam = f32c(lam).ravel()
        self._ck(self.lib.tdq_set_term_lambda(self.h, term, fptr(lam), lam.size))

    def get_lambda(self, term):
        out = np.empty(self.term_n[term], np.float32)
        self._ck(self.lib.tdq_get_term_lambda(self.h, term, fptr(out), out.size))
        return out

    def get_lambda_grad(self, term):
        out = np.empty(self.term_n[term], np.float32)
        self._ck(self.lib.tdq_get_term_lambda_grad(self.h, term, fptr(out), out.size))
        return out

    def get_residual(self, term):
        out = np.empty(self.term_n[term], np.float32)
        self._ck(self.lib.tdq_get_term_residual(self.h, term, fptr(out), out.size))
        return out

    # ---- hot path
    def loss_grad(self, n_vars=0):
        loss = C.c_float(0)
        g = np.empty(self.n_params, np.float32)
        vg = np.zeros(max(n_vars, 1), np.float32)
        self._ck(self.lib.tdq_loss_grad(self.h, C.byref(loss), fptr(g), fptr(vg)))
        return float(loss.value), g, vg[:n_vars]

    @staticmethod
    def adam_cfg(lr=0.005, beta1=0.99, beta2=0.999, eps=1e-7, sign=1):
        return tdq_adam_cfg(lr, beta1, beta2, eps, sign)

    def adam_steps(self, n_steps, w=None, lam=None, vars=None, want_hist=True):
        hist = np.zeros(max(n_steps, 1), np.float32)
        self._ck(self.lib.tdq_adam_steps(self.h, C.byref(w) if w is not None else None,
                                         C.byref(lam) if lam is not None else None,
                                         C.byref(vars) if vars is not None else None, int(n_steps),
                                         fptr(hist) if want_hist else None))
        return hist[:n_steps]

    def adam_epochs(self, n_epochs, terms, batch_sz, n_batches, w=None, lam=None, vars=None):
        """n_epochs minibatch epochs over consecutive slices of the listed terms' stored points, enqueued as one call"""
        hist = np.zeros(max(n_epochs, 1), np.float32)
        ids = (C.c_int32 * len(terms))(*[int(t) for t in terms])
        self._ck(self.lib.tdq_adam_epochs(self.h, C.byref(w) if w is not None else None,
                                          C.byref(lam) if lam is not None else None,
                                          C.byref(vars) if vars is not None else None, ids, len(terms), int(batch_sz),
                                          int(n_batches), int(n_epochs), fptr(hist)))
        return hist[:n_epochs]

    def adam_reset(self):
        self._ck(self.lib.tdq_adam_reset(self.h))

    def lbfgs(self, max_iter, lr=0.8):
        hist = np.zeros(max_iter + 2, np.float32)
        nh = C.c_int32(0)
        ne = C.c_int32(0)
        self._ck(self.lib.tdq_lbfgs(self.h, int(max_iter), float(lr), fptr(hist), C.byref(nh), C.byref(ne)))
        return hist[:nh.value].copy(), ne.value

    def lbfgs_hz(self, max_iter, tolerance=1e-20):
        """L-BFGS with the Hager-Zhang line search -> (f history, evaluations, status)"""
        hist = np.zeros(max_iter + 2, np.float32)
        nh, ne, st = C.c_int32(0), C.c_int32(0), C.c_int32(0)
        self._ck(self.lib.tdq_lbfgs_hz(self.h, int(max_iter), float(tolerance), fptr(hist), C.byref(nh), C.byref(ne), C.byref(st)))
        return hist[:nh.value], ne.value, st.value

    def predict(self, n_dims, orders, ops, out_slot, f_slot, X):
        X = f32c(X)
        spec, keep = make_spec(n_dims, orders, ops, out_slot, f_slot, 0)
        m = X.shape[0]
        u = np.empty(m, np.float32)
        f = np.zeros(m, np.float32)
        self._ck(self.lib.tdq_predict(self.h, C.byref(spec), fptr(X), m, fptr(u), fptr(f)))
        return u, f

    # ---- multi-GPU
    def unique_id(self):
        buf = C.create_string_buffer(128)
        if self.lib.tdq_comm_unique_id(buf):
            raise TdqError(self.lib.tdq_last_error(None).decode())
        return buf.raw

    def comm_init(self, uid, rank, world):
        buf = C.create_string_buffer(bytes(uid), 128)
        self._ck(self.lib.tdq_comm_init(self.h, buf, int(rank), int(world)))

    # ---- timing / info
    def comm_mode(self):
        buf = C.create_string_buffer(128)
        self._ck(self.lib.tdq_comm_mode(self.h, buf, 128))
        return buf.value.decode()

    def mark_begin(self):
        self._ck(self.lib.tdq_mark_begin(self.h))

    def mark_end(self):
        ms = C.c_float(0)
        self._ck(self.lib.tdq_mark_end(self.h, C.byref(ms)))
        return float(ms.value)

    def sync(self):
        self._ck(self.lib.tdq_sync(self.h))

    def set_kernel_timing(self, on):
        self._ck(self.lib.tdq_set_kernel_timing(self.h, 1 if on else 0))

    def term_kernel(self, term):
        buf = C.create_string_buffer(96)
        self._ck(self.lib.tdq_term_kernel(self.h, term, buf, 96))
        return buf.value.decode()

    def kernel_time(self, term, reset=True):
        ms = C.c_float(0)
        n = C.c_int64(0)
        self._ck(self.lib.tdq_kernel_time(self.h, term, 1 if reset else 0, C.byref(ms), C.byref(n)))
        return float(ms.value), int(n.value)

    def info(self):
        sm = C.c_int32(0)
        nl = C.c_int64(0)
        sb = C.c_int64(0)
        self._ck(self.lib.tdq_info(self.h, C.byref(sm), C.byref(nl), C.byref(sb)))
        return {"n_sm": sm.value, "launches": nl.value, "scratch_bytes": sb.value}
