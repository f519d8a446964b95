"""ctypes mirror of include/lsmc_b200.h — the C-ABI structs exactly as a P/Invoke host would lay them out.

Nothing here computes; it only owns numpy buffers and hands their addresses across the boundary.
"""
import ctypes as C
import numpy as np

LSMC_OK, LSMC_EINVAL, LSMC_ECONSTRAINT, LSMC_ECANCELLED, LSMC_ECUDA, LSMC_EUNSUPPORTED = range(6)
CONSTRAINT_CONSTANT, CONSTRAINT_PIECEWISE_LINEAR, CONSTRAINT_STEP, CONSTRAINT_POLYNOMIAL = 0, 1, 2, 3
RNG_PHILOX, RNG_NORMALS, RNG_PANELS, RNG_MT19937, RNG_MT19937_CONTINUE = 0, 1, 2, 3, 4
MAX_FACTORS, MAX_BASIS, NUM_TRIGGER_VOLUMES = 8, 64, 10

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class LsmcProblem(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("n_periods", C.c_int32), ("n_factors", C.c_int32), ("n_basis", C.c_int32),
        ("n_sims", C.c_int64), ("n_sims_global", C.c_int64), ("sim_offset", C.c_int64),
        ("extra_decisions", C.c_int32), ("first_period_is_current", C.c_int32),
        ("discount_deltas", C.c_int32), ("must_be_empty_at_end", C.c_int32),
        ("numerical_tolerance", C.c_double), ("start_inventory", C.c_double), ("grid_spacing", C.c_double),
        ("terminal_coef", C.c_double * 4),
        ("sim_time", _dp), ("mean_reversion", _dp), ("factor_vol", _dp), ("factor_corr", _dp), ("forward", _dp),
        ("spot_power", _ip), ("factor_power", _ip),
        ("min_inventory", _dp), ("max_inventory", _dp), ("constraint_id", _ip),
        ("inject_cost", _dp), ("inject_cost_df", _dp), ("withdraw_cost", _dp), ("withdraw_cost_df", _dp),
        ("consumed_inject", _dp), ("consumed_withdraw", _dp), ("inventory_loss", _dp),
        ("inventory_cost", _dp), ("inventory_cost_df", _dp), ("settle_df", _dp),
        ("n_constraints", C.c_int32), ("n_pillars", C.c_int32),
        ("constraint_kind", _ip), ("constraint_first", _ip), ("constraint_count", _ip),
        ("pillar_inventory", _dp), ("pillar_min_rate", _dp), ("pillar_max_rate", _dp),
    ]


class LsmcRng(C.Structure):
    _fields_ = [("mode", C.c_int32), ("antithetic", C.c_int32), ("seed", C.c_uint64),
                ("normals", _dp), ("spot", _dp), ("factors", _dp)]


class LsmcLayout(C.Structure):
    _fields_ = [("n_periods", C.c_int32), ("n_sim_periods", C.c_int32), ("max_grid", C.c_int32),
                ("max_decisions", C.c_int32), ("n_sims", C.c_int64)]


class LsmcResults(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("reserved", C.c_int32),
        ("npv", C.c_double), ("npv_std_error", C.c_double), ("backward_npv", C.c_double),
        ("deltas", _dp), ("profile", _dp), ("trigger_summary", _dp), ("trigger_inject", _dp),
        ("trigger_withdraw", _dp), ("pv_by_sim", _dp),
        ("spot_regress", _dp), ("spot_valuation", _dp), ("inventory", _dp), ("inject_withdraw", _dp),
        ("consumed", _dp), ("loss", _dp), ("net_volume", _dp), ("pv", _dp),
        ("decision_index", C.POINTER(C.c_int8)),
    ]


class LsmcTimings(C.Structure):
    _fields_ = [("regression_sim_ms", C.c_double), ("backward_ms", C.c_double), ("valuation_sim_ms", C.c_double),
                ("forward_ms", C.c_double), ("total_ms", C.c_double), ("backward_sweep_kernel_ms", C.c_double),
                ("backward_solve_kernel_ms", C.c_double), ("backward_sweep_launches", C.c_int64),
                ("kernel_launches", C.c_int64), ("sweep_bytes", C.c_double), ("span_ms", C.c_double)]


PROGRESS_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_double)

PANEL_NAMES = ("spot_regress", "spot_valuation", "inventory", "inject_withdraw", "consumed", "loss", "net_volume", "pv")


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ptr(a, typ):
    return a.ctypes.data_as(typ)


class Problem:
    """Flattened LsmcValuationParameters: numpy arrays + scalars; `.c_struct()` gives the ABI view.

    Field meanings are documented in include/lsmc_b200.h.  Period arrays have P entries.
    """
    _F64 = ("sim_time", "mean_reversion", "factor_vol", "factor_corr", "forward", "min_inventory", "max_inventory",
            "inject_cost", "inject_cost_df", "withdraw_cost", "withdraw_cost_df", "consumed_inject",
            "consumed_withdraw", "inventory_loss", "inventory_cost", "inventory_cost_df", "settle_df",
            "pillar_inventory", "pillar_min_rate", "pillar_max_rate")
    _I32 = ("spot_power", "factor_power", "constraint_id", "constraint_kind", "constraint_first", "constraint_count")

    def __init__(self, **kw):
        self.n_sims_global = 0
        self.sim_offset = 0
        self.extra_decisions = 0
        self.first_period_is_current = 0
        self.discount_deltas = 0
        self.must_be_empty_at_end = 1
        self.numerical_tolerance = 1e-12
        self.terminal_coef = (0.0, 0.0, 0.0, 0.0)
        for k, v in kw.items():
            setattr(self, k, v)

    def finalize(self):
        for name in self._F64:
            setattr(self, name, _f64(getattr(self, name)))
        for name in self._I32:
            setattr(self, name, _i32(getattr(self, name)))
        P = int(self.n_periods)
        F = int(self.n_factors)
        K = int(self.n_basis)
        assert self.sim_time.shape == (P,) and self.forward.shape == (P,)
        assert self.factor_vol.shape == (P, F) and self.factor_corr.shape == (F, F)
        assert self.spot_power.shape == (K,) and self.factor_power.shape == (K, F)
        for name in ("min_inventory", "max_inventory", "constraint_id", "inject_cost", "inject_cost_df", "withdraw_cost",
                     "withdraw_cost_df", "consumed_inject", "consumed_withdraw", "inventory_loss", "inventory_cost",
                     "inventory_cost_df", "settle_df"):
            assert getattr(self, name).shape == (P,), name
        return self

    def copy(self, **overrides):
        import copy
        other = copy.copy(self)
        for k, v in overrides.items():
            setattr(other, k, v)
        return other

    @property
    def n_sim_periods(self):
        return int(self.n_periods) - int(self.first_period_is_current)

    def c_struct(self):
        self.finalize()
        s = LsmcProblem()
        s.struct_size = C.sizeof(LsmcProblem)
        s.n_periods, s.n_factors, s.n_basis = int(self.n_periods), int(self.n_factors), int(self.n_basis)
        s.n_sims, s.n_sims_global, s.sim_offset = int(self.n_sims), int(self.n_sims_global), int(self.sim_offset)
        s.extra_decisions = int(self.extra_decisions)
        s.first_period_is_current = int(self.first_period_is_current)
        s.discount_deltas = int(bool(self.discount_deltas))
        s.must_be_empty_at_end = int(bool(self.must_be_empty_at_end))
        s.numerical_tolerance = float(self.numerical_tolerance)
        s.start_inventory = float(self.start_inventory)
        s.grid_spacing = float(self.grid_spacing)
        s.terminal_coef = (C.c_double * 4)(*[float(x) for x in self.terminal_coef])
        for name in self._F64:
            setattr(s, name, _ptr(getattr(self, name), _dp))
        for name in self._I32:
            setattr(s, name, _ptr(getattr(self, name), _ip))
        s.n_constraints = int(self.constraint_kind.shape[0])
        s.n_pillars = int(self.pillar_inventory.shape[0])
        s._keepalive = self
        return s


class Rng:
    """One simulation set's draw source (lsmc_rng)."""

    def __init__(self, mode=RNG_PHILOX, seed=0, antithetic=True, normals=None, spot=None, factors=None):
        self.mode, self.seed, self.antithetic = mode, int(seed), bool(antithetic)
        self.normals = None if normals is None else _f64(normals)
        self.spot = None if spot is None else _f64(spot)
        self.factors = None if factors is None else _f64(factors)

    @staticmethod
    def philox(seed, antithetic=True):
        return Rng(RNG_PHILOX, seed, antithetic)

    @staticmethod
    def mt19937(seed, antithetic=True):
        """The reference's seeded stream: MersenneTwisterGenerator(seed, antithetic) of LsmcValuationParameters.cs:181-191."""
        return Rng(RNG_MT19937, int(seed) & 0xFFFFFFFF, antithetic)

    @staticmethod
    def mt19937_continue(antithetic=True):
        """The generator of the context's previous MT19937 simulation continues (LVP.cs:186-187: valuationSimSeed == null)."""
        return Rng(RNG_MT19937_CONTINUE, 0, antithetic)

    @staticmethod
    def from_normals(normals):
        return Rng(RNG_NORMALS, normals=normals)

    @staticmethod
    def from_panels(spot, factors):
        return Rng(RNG_PANELS, spot=spot, factors=factors)

    def c_struct(self):
        s = LsmcRng()
        s.mode, s.antithetic, s.seed = self.mode, int(self.antithetic), self.seed & 0xFFFFFFFFFFFFFFFF
        if self.normals is not None:
            s.normals = _ptr(self.normals, _dp)
        if self.spot is not None:
            s.spot = _ptr(self.spot, _dp)
        if self.factors is not None:
            s.factors = _ptr(self.factors, _dp)
        s._keepalive = self
        return s


class Results:
    """Caller-allocated result buffers (lsmc_results)."""

    def __init__(self, P, S, N, panels=(), pv_by_sim=True, decision_index=False, triggers=True, pinned=False):
        """pinned: allocate the per-sim panels in page-locked host memory (lsmc_host_alloc) so that the library's copy
        engine writes them directly."""
        self.P, self.S, self.N = int(P), int(S), int(N)
        if pinned:
            from ._native import pinned_zeros
            panel_zeros = pinned_zeros
        else:
            panel_zeros = np.zeros
        self.deltas = np.zeros(P)
        self.profile = np.zeros((P, 6))
        self.trigger_summary = np.full((P - 1, 4), np.nan) if triggers else None
        self.trigger_inject = np.full((P - 1, NUM_TRIGGER_VOLUMES, 2), np.nan) if triggers else None
        self.trigger_withdraw = np.full((P - 1, NUM_TRIGGER_VOLUMES, 2), np.nan) if triggers else None
        self.pv_by_sim = np.zeros(N) if pv_by_sim else None
        for name in PANEL_NAMES:
            rows = S if name.startswith("spot_") else P
            setattr(self, name, panel_zeros((rows, N)) if (panels == "all" or name in panels) else None)
        self.decision_index = np.zeros((P - 1, N), dtype=np.int8) if decision_index else None
        self.npv = self.npv_std_error = self.backward_npv = float("nan")
        self._c = None

    def c_struct(self):
        s = LsmcResults()
        s.struct_size = C.sizeof(LsmcResults)
        for name in ("deltas", "profile", "trigger_summary", "trigger_inject", "trigger_withdraw", "pv_by_sim") + PANEL_NAMES:
            a = getattr(self, name)
            if a is not None:
                setattr(s, name, _ptr(a, _dp))
        if self.decision_index is not None:
            s.decision_index = self.decision_index.ctypes.data_as(C.POINTER(C.c_int8))
        self._c = s
        return s

    def sync_scalars(self):
        self.npv, self.npv_std_error, self.backward_npv = self._c.npv, self._c.npv_std_error, self._c.backward_npv
        return self
