"""ctypes wrapper of the CPU oracle (oracle/liblsmc_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under storage_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from storage_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liblsmc_oracle.so")
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class OracleExtras(C.Structure):
    _fields_ = [("max_grid", C.c_int32), ("reserved", C.c_int32),
                ("inventory_space_min", _dp), ("inventory_space_max", _dp), ("grid_count", _ip), ("grids", _dp),
                ("regression_coeffs", _dp), ("values", _dp), ("factors_regress", _dp), ("factors_valuation", _dp),
                ("current_continuation", _dp), ("timing_ms", C.c_double * 5)]


class OracleError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("oracle status %d: %s" % (status, message))
        self.status = status


def build(force=False):
    """Compile the oracle with the committed Makefile (g++ only; the C# reference cannot be built here)."""
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "lsmc_oracle.cpp")):
        subprocess.check_call(["make", "-C", _HERE, "liblsmc_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.oracle_last_error.restype = C.c_char_p
        _lib.oracle_det_exp.restype = C.c_double
        _lib.oracle_det_exp.argtypes = [C.c_double]
    return _lib


def _check(status):
    if status != 0:
        raise OracleError(status, lib().oracle_last_error().decode())


def _p(a):
    return a.ctypes.data_as(_dp)


class OracleOutput:
    pass


def calculate(problem, regression_rng, valuation_rng, panels=(), keep_values=False, keep_factors=False,
              decision_index=True, triggers=True, max_grid=None):
    """Run the oracle end to end.  Returns an object with `.results` (an _abi.Results) and the extras."""
    P, S, N, K, F = problem.n_periods, problem.n_sim_periods, int(problem.n_sims), problem.n_basis, problem.n_factors
    res = _abi.Results(P, S, N, panels=panels, decision_index=decision_index, triggers=triggers)
    if max_grid is None:
        span = float(np.max(problem.max_inventory) - np.min(problem.min_inventory))
        max_grid = int(np.ceil(span / problem.grid_spacing)) + 4
    ex = OracleExtras()
    ex.max_grid = max_grid
    out = OracleOutput()
    out.inventory_space_min = np.zeros(P)
    out.inventory_space_max = np.zeros(P)
    out.grid_count = np.zeros(P, dtype=np.int32)
    out.grids = np.zeros((P, max_grid))
    out.regression_coeffs = np.zeros((P - 1, max_grid, K))
    out.current_continuation = np.zeros(max_grid)
    ex.inventory_space_min, ex.inventory_space_max = _p(out.inventory_space_min), _p(out.inventory_space_max)
    ex.grid_count = out.grid_count.ctypes.data_as(_ip)
    ex.grids, ex.regression_coeffs = _p(out.grids), _p(out.regression_coeffs)
    ex.current_continuation = _p(out.current_continuation)
    out.values = None
    if keep_values:
        out.values = np.zeros((P, max_grid, N))
        ex.values = _p(out.values)
    out.factors_regress = out.factors_valuation = None
    if keep_factors:
        out.factors_regress = np.zeros((S, F, N))
        out.factors_valuation = np.zeros((S, F, N))
        ex.factors_regress, ex.factors_valuation = _p(out.factors_regress), _p(out.factors_valuation)
    cp, cr, cv, cres = problem.c_struct(), regression_rng.c_struct(), valuation_rng.c_struct(), res.c_struct()
    _check(lib().oracle_lsmc_calculate(C.byref(cp), C.byref(cr), C.byref(cv), C.byref(cres), C.byref(ex)))
    res.sync_scalars()
    out.results = res
    out.timing_ms = dict(zip(("regression_sim", "backward", "valuation_sim", "forward", "total"), list(ex.timing_ms)))
    return out


def num_threads():
    return lib().oracle_num_threads()


def set_num_threads(n):
    lib().oracle_set_num_threads(int(n))


def bang_bang_decision_set(rmin, rmax, inventory, loss, next_min, next_max, tol, extra=0):
    out = np.zeros(2 * extra + 3)
    count = C.c_int()
    _check(lib().oracle_bang_bang_decision_set(C.c_double(rmin), C.c_double(rmax), C.c_double(inventory), C.c_double(loss),
                                               C.c_double(next_min), C.c_double(next_max), C.c_double(tol), C.c_int(extra),
                                               _p(out), C.byref(count)))
    return out[:count.value].copy()


def bisect_inventory_space(grid, inventory, tol):
    g = np.ascontiguousarray(grid, dtype=np.float64)
    lo, hi = C.c_int(), C.c_int()
    _check(lib().oracle_bisect_inventory_space(_p(g), C.c_int(len(g)), C.c_double(inventory), C.c_double(tol),
                                               C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def max_value_and_index(values):
    a = np.ascontiguousarray(values, dtype=np.float64)
    mx, idx = C.c_double(), C.c_int()
    _check(lib().oracle_max_value_and_index(_p(a), C.c_int(len(a)), C.byref(mx), C.byref(idx)))
    return mx.value, idx.value


def grid_points(lo, hi, spacing, cap=100000):
    out = np.zeros(cap)
    count = C.c_int()
    _check(lib().oracle_grid_points(C.c_double(lo), C.c_double(hi), C.c_double(spacing), _p(out), C.c_int(cap), C.byref(count)))
    return out[:count.value].copy()


def inventory_space(problem):
    P = problem.n_periods
    lo, hi = np.zeros(P - 1), np.zeros(P - 1)
    cp = problem.c_struct()
    _check(lib().oracle_inventory_space(C.byref(cp), _p(lo), _p(hi)))
    return lo, hi


def inject_withdraw_range(problem, period, inventory):
    rmin, rmax = C.c_double(), C.c_double()
    cp = problem.c_struct()
    _check(lib().oracle_inject_withdraw_range(C.byref(cp), C.c_int(period), C.c_double(inventory), C.byref(rmin), C.byref(rmax)))
    return rmin.value, rmax.value


def constraint_bounds(problem, period, next_lo, next_hi):
    up, lo = C.c_double(), C.c_double()
    cp = problem.c_struct()
    _check(lib().oracle_constraint_bounds(C.byref(cp), C.c_int(period), C.c_double(next_lo), C.c_double(next_hi),
                                          C.byref(up), C.byref(lo)))
    return up.value, lo.value


def design_matrix(problem, spot, factors):
    spot = np.ascontiguousarray(spot, dtype=np.float64)
    factors = np.ascontiguousarray(factors, dtype=np.float64)
    N = spot.shape[0]
    X = np.zeros((problem.n_basis, N))
    cp = problem.c_struct()
    _check(lib().oracle_design_matrix(C.byref(cp), _p(spot), _p(factors), C.c_int64(N), _p(X)))
    return X.T.copy()   # N x K


def simulate(problem, rng):
    S, F, N = problem.n_sim_periods, problem.n_factors, int(problem.n_sims)
    spot, fac = np.zeros((S, N)), np.zeros((S, F, N))
    cp, cr = problem.c_struct(), rng.c_struct()
    _check(lib().oracle_simulate(C.byref(cp), C.byref(cr), _p(spot), _p(fac)))
    return spot, fac


def regress(X, y):
    X = np.asarray(X, dtype=np.float64)
    N, K = X.shape
    Xc = np.ascontiguousarray(X.T)   # column-major N x K == row-major K x N
    y = np.ascontiguousarray(y, dtype=np.float64)
    beta, fitted = np.zeros(K), np.zeros(N)
    _check(lib().oracle_regress(_p(Xc), C.c_int64(N), C.c_int(K), _p(y), _p(beta), _p(fitted)))
    return beta, fitted


def det_exp(x):
    return lib().oracle_det_exp(float(x))


def philox4x32_10(ctr, key):
    c = (C.c_uint32 * 4)(*[int(x) for x in ctr])
    lib().oracle_philox4x32_10(c, C.c_uint32(int(key[0])), C.c_uint32(int(key[1])))
    return [int(x) for x in c]


def philox_normal_pair(seed, draw, step, pair):
    z0, z1 = C.c_double(), C.c_double()
    lib().oracle_philox_normal_pair(C.c_uint64(seed), C.c_uint64(draw), C.c_uint32(step), C.c_uint32(pair), C.byref(z0), C.byref(z1))
    return z0.value, z1.value
