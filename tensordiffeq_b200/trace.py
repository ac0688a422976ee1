"""Tracing of the user's residual callbacks into tdq_op programs.

The reference runs `f_model(u_model, *cols)` (models.py:157) and
`deriv_model(u_model, *cols)` (boundaries.py:238-239) under tf.function: the
callback builds the PDE residual out of `u_model(tf.concat([...], 1))`,
`tf.gradients`, `tf.constant`, arithmetic and tf.math functions
(e.g. examples/AC-SA.py:36-44, examples/CH-discovery.py:20-35).

Here the callback is called ONCE with symbolic arguments.  `tf.gradients(y, x)`
becomes a symbolic total derivative: d/dx_i of the jet leaf D^alpha u is the leaf
D^(alpha+e_i) u, everything else follows the chain rule.  The result is an
expression over jet leaves, coordinates, constants, trainable coefficients, the
per-point target and the per-point lambda, which is lowered to the straight-line
program the CUDA kernel interprets per collocation point (include/tdq_b200.h).
"""
import math

import numpy as np

F32 = np.float32
_UNARY = {"neg", "sin", "cos", "exp", "tanh", "sqrt", "log"}
_BINARY = {"add", "sub", "mul", "div"}


class Sym:
    """A node of the expression DAG.  `listy` remembers that a Python list took
    part (tf.gradients returns a list; the reference then carries a leading axis
    of 1 through the residual -- SURVEY quirk Q1 -- which predict() reproduces)."""
    __slots__ = ("kind", "args", "val", "listy", "_key")
    __array_priority__ = 10000
    __array_ufunc__ = None

    _intern = {}

    def __new__(cls, kind, args=(), val=None, listy=False):
        # hash-consing: structurally equal nodes are one object, so common sub-expressions
        # (plentiful after symbolic differentiation) are found by identity
        args = tuple(args)
        listy = bool(listy or any(a.listy for a in args))
        if isinstance(val, np.floating):
            vkey = ("f", float(val))
        else:
            vkey = val
        k = (kind, tuple(id(a) for a in args), vkey, listy)
        node = cls._intern.get(k)
        if node is None:
            node = object.__new__(cls)
            node.kind = kind
            node.args = args
            node.val = val
            node.listy = listy
            node._key = k[:3]      # the list flag only affects the reported shape, not the program
            cls._intern[k] = node
        return node

    def key(self):
        return self._key

    def is_const(self, v=None):
        return self.kind == "const" and (v is None or self.val == F32(v))

    # -- operators (python scalars, numpy scalars, 1-element lists accepted)
    def __add__(self, o): return add(self, o)
    def __radd__(self, o): return add(o, self)
    def __sub__(self, o): return sub(self, o)
    def __rsub__(self, o): return sub(o, self)
    def __mul__(self, o): return mul(self, o)
    def __rmul__(self, o): return mul(o, self)
    def __truediv__(self, o): return div(self, o)
    def __rtruediv__(self, o): return div(o, self)
    def __neg__(self): return neg(self)
    def __pow__(self, o): return power(self, o)

    def __getitem__(self, idx):
        # tf.gradients(...)[0] on something that already is an expression, u[:, 0:1] etc.
        return self

    def numpy(self):
        if self.kind == "const":
            return np.float32(self.val)
        raise TypeError("symbolic value has no numpy() during tracing")

    def __repr__(self):
        return "Sym(%s%s)" % (self.kind, "" if self.val is None else ":%r" % (self.val,))


def const(v):
    return Sym("const", (), F32(v))


def as_sym(x):
    if isinstance(x, Sym):
        return x
    if isinstance(x, (list, tuple)):
        if len(x) != 1:
            raise TypeError("only single-element lists (tf.gradients results) take part in residual arithmetic")
        s = as_sym(x[0])
        return Sym(s.kind, s.args, s.val, listy=True)
    if hasattr(x, "sym_value"):
        return x.sym_value()
    if isinstance(x, (int, float, np.floating, np.integer)):
        return const(x)
    if isinstance(x, np.ndarray) and x.size == 1:
        return const(x.reshape(-1)[0])
    if hasattr(x, "numpy"):
        v = np.asarray(x.numpy())
        if v.size == 1:
            return const(v.reshape(-1)[0])
    raise TypeError("cannot use %r inside a traced residual" % (type(x),))


def _fold2(kind, a, b):
    x, y = F32(a.val), F32(b.val)
    with np.errstate(all="ignore"):
        r = {"add": x + y, "sub": x - y, "mul": x * y, "div": x / y}[kind]
    return Sym("const", (), F32(r), listy=a.listy or b.listy)


def add(a, b):
    a, b = as_sym(a), as_sym(b)
    if a.is_const() and b.is_const(): return _fold2("add", a, b)
    if a.is_const(0): return b if not a.listy else Sym(b.kind, b.args, b.val, True)
    if b.is_const(0): return a if not b.listy else Sym(a.kind, a.args, a.val, True)
    return Sym("add", (a, b))


def sub(a, b):
    a, b = as_sym(a), as_sym(b)
    if a.is_const() and b.is_const(): return _fold2("sub", a, b)
    if b.is_const(0): return a if not b.listy else Sym(a.kind, a.args, a.val, True)
    if a.is_const(0): return neg(b)
    return Sym("sub", (a, b))


def mul(a, b):
    a, b = as_sym(a), as_sym(b)
    if a.is_const() and b.is_const(): return _fold2("mul", a, b)
    if a.is_const(0) or b.is_const(0): return Sym("const", (), F32(0), listy=a.listy or b.listy)
    if a.is_const(1): return b if not a.listy else Sym(b.kind, b.args, b.val, True)
    if b.is_const(1): return a if not b.listy else Sym(a.kind, a.args, a.val, True)
    return Sym("mul", (a, b))


def div(a, b):
    a, b = as_sym(a), as_sym(b)
    if a.is_const() and b.is_const(): return _fold2("div", a, b)
    if a.is_const(0): return Sym("const", (), F32(0), listy=a.listy or b.listy)
    if b.is_const(1): return a
    return Sym("div", (a, b))


def neg(a):
    a = as_sym(a)
    if a.is_const(): return Sym("const", (), F32(-a.val), listy=a.listy)
    if a.kind == "neg": return a.args[0]
    return Sym("neg", (a,))


def power(a, n):
    a = as_sym(a)
    if isinstance(n, Sym):
        if not n.is_const():
            raise TypeError("only constant exponents are supported")
        n = float(n.val)
    n = float(n)
    if a.is_const():
        return Sym("const", (), F32(np.power(F32(a.val), F32(n))), listy=a.listy)
    if n == 1.0: return a
    if n == 0.0: return const(1.0)
    if n == int(n) and abs(n) <= 16:
        return Sym("powi", (a,), int(n))
    return Sym("powf", (a,), F32(n))


def unary(kind, a):
    a = as_sym(a)
    if a.is_const():
        with np.errstate(all="ignore"):
            f = {"sin": np.sin, "cos": np.cos, "exp": np.exp, "tanh": np.tanh, "sqrt": np.sqrt, "log": np.log}[kind]
            return Sym("const", (), F32(f(F32(a.val))), listy=a.listy)
    return Sym(kind, (a,))


def jet(alpha):
    return Sym("jet", (), tuple(int(v) for v in alpha))


def coord(i):
    return Sym("coord", (), int(i))


def var(i):
    return Sym("var", (), int(i))


TARGET = Sym("target")
LAMBDA = Sym("lam")


class MixedPartialError(Exception):
    pass


def total_derivative(e, i, memo=None):
    """d e / d x_i along the network input x_i (what tf.gradients(e, x_i)[0] evaluates to)."""
    if memo is None:
        memo = {}
    k = id(e)
    if k in memo:
        return memo[k]
    kd = e.kind
    if kd == "jet":
        a = list(e.val)
        a[i] += 1
        if a[i] > 4:
            raise NotImplementedError("derivatives above 4th order are not supported by the jet kernels")
        if sum(1 for v in a if v) > 1 and not (len(a) == 2 and tuple(a) == (1, 1)):
            # the jets carry pure derivatives per input direction, plus u_xy for two-dimensional second-order problems
            raise MixedPartialError("mixed partial derivatives (%r) are not supported by the jet kernels "
                                    "(only u_xy of a two-dimensional problem is)" % (tuple(a),))
        r = jet(a)
    elif kd == "coord":
        r = const(1.0 if e.val == i else 0.0)
    elif kd in ("const", "var", "target", "lam"):
        r = const(0.0)
    elif kd == "add":
        r = add(total_derivative(e.args[0], i, memo), total_derivative(e.args[1], i, memo))
    elif kd == "sub":
        r = sub(total_derivative(e.args[0], i, memo), total_derivative(e.args[1], i, memo))
    elif kd == "mul":
        x, y = e.args
        r = add(mul(total_derivative(x, i, memo), y), mul(x, total_derivative(y, i, memo)))
    elif kd == "div":
        x, y = e.args
        r = sub(div(total_derivative(x, i, memo), y), mul(div(e, y), total_derivative(y, i, memo)))
    elif kd == "neg":
        r = neg(total_derivative(e.args[0], i, memo))
    elif kd == "sin":
        r = mul(unary("cos", e.args[0]), total_derivative(e.args[0], i, memo))
    elif kd == "cos":
        r = neg(mul(unary("sin", e.args[0]), total_derivative(e.args[0], i, memo)))
    elif kd == "exp":
        r = mul(e, total_derivative(e.args[0], i, memo))
    elif kd == "tanh":
        r = mul(sub(const(1.0), mul(e, e)), total_derivative(e.args[0], i, memo))
    elif kd == "sqrt":
        r = div(total_derivative(e.args[0], i, memo), mul(const(2.0), e))
    elif kd == "log":
        r = div(total_derivative(e.args[0], i, memo), e.args[0])
    elif kd == "powi":
        n = e.val
        r = mul(mul(const(float(n)), power(e.args[0], n - 1)), total_derivative(e.args[0], i, memo))
    elif kd == "powf":
        n = float(e.val)
        r = mul(mul(const(n), power(e.args[0], n - 1.0)), total_derivative(e.args[0], i, memo))
    else:
        raise TypeError(kd)
    memo[k] = r
    return r


class Orders(list):
    """highest pure derivative per input dimension; .mixed = 1 when the mixed partial u_xy is read as well (it then is
    the jet channel after the pure ones, tdq_term_spec.mixed)"""
    mixed = 0


def jet_orders(exprs, n_dims):
    """highest pure derivative per input dimension among the jet leaves of the expressions"""
    orders = Orders([0] * n_dims)
    seen = set()

    def walk(e):
        if id(e) in seen:
            return
        seen.add(id(e))
        if e.kind == "jet":
            for i, v in enumerate(e.val):
                orders[i] = max(orders[i], v)
            if sum(1 for v in e.val if v) > 1:
                orders.mixed = 1
        for a in e.args:
            walk(a)

    for e in exprs:
        walk(e)
    if orders.mixed and max(orders) > 2:
        raise MixedPartialError("the mixed partial u_xy is carried next to pure derivatives of order <= 2 only")
    return orders


def lower(outputs, n_dims, orders=None):
    """Lower expressions to one straight-line program with common sub-expressions shared.
    outputs: dict name -> Sym.  Returns (ops, slots {name: slot}, orders)."""
    if orders is None:
        orders = jet_orders(outputs.values(), n_dims)
    offs = [1]
    for i in range(n_dims):
        offs.append(offs[-1] + orders[i])
    ops, index = [], {}

    def chan(alpha):
        nz = [(i, v) for i, v in enumerate(alpha) if v]
        if not nz:
            return 0
        if len(nz) > 1:
            return offs[-1]                  # the mixed partial: the channel after the pure ones
        (i, v), = nz
        return offs[i] + v - 1

    def emit(e):
        k = e.key()
        if k in index:
            return index[k]
        kd = e.kind
        if kd == "const":
            op = ("CONST", 0, 0, float(e.val))
        elif kd == "jet":
            op = ("JET", chan(e.val), 0, 0.0)
        elif kd == "coord":
            op = ("COORD", e.val, 0, 0.0)
        elif kd == "var":
            op = ("VAR", e.val, 0, 0.0)
        elif kd == "target":
            op = ("TARGET", 0, 0, 0.0)
        elif kd == "lam":
            op = ("LAMBDA", 0, 0, 0.0)
        elif kd in _BINARY:
            a, b = emit(e.args[0]), emit(e.args[1])
            op = (kd.upper(), a, b, 0.0)
        elif kd in _UNARY:
            op = (kd.upper(), emit(e.args[0]), 0, 0.0)
        elif kd == "powi":
            op = ("POWI", emit(e.args[0]), 0, float(e.val))
        elif kd == "powf":
            op = ("POWF", emit(e.args[0]), 0, float(e.val))
        else:
            raise TypeError(kd)
        ops.append(op)
        index[k] = len(ops) - 1
        return index[k]

    slots = {name: emit(e) for name, e in outputs.items()}
    return ops, slots, orders


# ---------------------------------------------------------------------------
# numpy evaluation of a lowered program (host-side checks / CPU tests of the tracer;
# NOT a compute path of the product)
# ---------------------------------------------------------------------------
def eval_program(ops, jets, coords, target=None, lam=None, vars_=None):
    """jets: [C][n], coords: [d][n] -> list of slot arrays"""
    slot = []
    for (op, a, b, imm) in ops:
        if op == "CONST": v = np.full_like(coords[0], imm, dtype=np.float64)
        elif op == "JET": v = jets[a]
        elif op == "COORD": v = coords[a]
        elif op == "TARGET": v = target
        elif op == "LAMBDA": v = lam
        elif op == "VAR": v = np.full_like(coords[0], vars_[a], dtype=np.float64)
        elif op == "ADD": v = slot[a] + slot[b]
        elif op == "SUB": v = slot[a] - slot[b]
        elif op == "MUL": v = slot[a] * slot[b]
        elif op == "DIV": v = slot[a] / slot[b]
        elif op == "NEG": v = -slot[a]
        elif op == "SIN": v = np.sin(slot[a])
        elif op == "COS": v = np.cos(slot[a])
        elif op == "EXP": v = np.exp(slot[a])
        elif op == "TANH": v = np.tanh(slot[a])
        elif op == "SQRT": v = np.sqrt(slot[a])
        elif op == "LOG": v = np.log(slot[a])
        elif op == "POWI": v = slot[a] ** int(imm)
        elif op == "POWF": v = slot[a] ** imm
        else: raise ValueError(op)
        slot.append(v)
    return slot
