"""Import placeholder (plotting is out of scope)."""
import types as _t, sys as _s
for _n in ("pyplot", "gridspec"):
    _m = _t.ModuleType("matplotlib." + _n)
    _s.modules["matplotlib." + _n] = _m
    globals()[_n] = _m
def use(*a, **k):
    pass
rcParams = {}
