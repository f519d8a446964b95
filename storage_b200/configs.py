"""Synthetic valuation configs of BASELINE.json: built through the public host API
(CmdtyStorage + flatten_problem), so benchmark and parity inputs exercise the same flattening as user calls."""
import numpy as np
import pandas as pd

from .cmdty_storage import CmdtyStorage, RatchetInterp
from .flatten import flatten_problem
from .multi_factor_params import create_3_factor_season_params

README_BASIS = '1 + x_st + x_sw + x_lt + s + x_st**2 + x_sw**2 + x_lt**2 + s**2 + s * x_st'


def _settle_20th_next_month(period):   # README.md:133-134, LSMCTest:152
    return (period.asfreq('M') + 1).asfreq('D', 's') + 19


def _sin_forward(index, val_period):   # LsmcBenchmarks.cs:78-84
    days = np.array([(p.asfreq('D', 's') - val_period.asfreq('D', 's')).n for p in index], dtype=np.float64)
    return pd.Series(53.5 + 24.6 * np.sin(2.0 * np.pi / 365.0 * days), index=index)


def _seasonal_basis(basis):
    return basis.replace('x_st', 'x0').replace('x_lt', 'x1').replace('x_sw', 'x2')


def _ratchet_storage(start, end, years=1, **kw):   # README.md:100-125 shape
    ratchets = []
    y0 = pd.Period(start, freq='D')
    for y in range(years):
        a = y0 + 365 * y
        ratchets += [
            (a, [(0.0, -150.0, 250.0), (2000.0, -200.0, 175.0), (5000.0, -260.0, 155.0), (7000.0, -275.0, 132.0)]),
            (a + 153, [(0.0, -130.0, 260.0), (2000.0, -190.0, 190.0), (5000.0, -230.0, 165.0), (7000.0, -245.0, 148.0)]),
        ]
    return CmdtyStorage('D', start, end, 0.1, 0.25, ratchets=ratchets, ratchet_interp=RatchetInterp.LINEAR, **kw)


def build_config(name, num_sims=None, start_inventory=None, basis=None, end=None, storage=None, grid_points=None):
    """Returns a flattened Problem for config 'c1'..'c5' (BASELINE.json configs[0..4]).

    `basis`, `end`, `storage` and `grid_points` override the config's basis-function string, storage end date, storage object
    and number of inventory grid points (parity-test variants of the named shapes: other K, shorter horizons, other constraint
    kinds, grids beyond the pipelined sweep's 104 rows)."""
    name = name.lower()
    basis_override, end_override, storage_override = basis, end, storage
    if name == 'c1':    # README simple storage, three-factor seasonal, 1k paths
        N = num_sims or 1000
        start, end, val = '2021-04-01', '2022-04-01', '2021-04-01'
        st = CmdtyStorage('D', start, end, 0.01, 0.025, min_inventory=0.0, max_inventory=1400.0,
                          max_injection_rate=25.5, max_withdrawal_rate=30.9)
        factors_args = dict(seasonal=(91.0, 0.85, 0.30, 0.19)); basis = _seasonal_basis(README_BASIS); inv = 0.0
    elif name == 'c2':  # one-factor, constant rates, 100k paths
        N = num_sims or 100_000
        start, end, val = '2021-04-01', '2022-04-01', '2021-04-01'
        st = CmdtyStorage('D', start, end, 0.01, 0.025, min_inventory=0.0, max_inventory=1500.0,
                          max_injection_rate=25.5, max_withdrawal_rate=30.9)
        factors_args = dict(one_factor=(12.5, 0.95)); basis = '1 + x0 + x0**2'; inv = 0.0
    elif name == 'c3':  # three-factor seasonal, ratchets + losses, 1M paths
        N = num_sims or 1_000_000
        start, end, val = '2021-04-01', '2022-04-01', '2021-04-01'
        st = _ratchet_storage(start, end, inventory_loss=0.0005, cmdty_consumed_inject=0.01)
        factors_args = dict(seasonal=(91.0, 0.85, 0.30, 0.19)); basis = _seasonal_basis(README_BASIS); inv = 0.0
    elif name == 'c4':  # five-factor, degree-3 polynomial basis, two years
        N = num_sims or 4_000_000
        start, end, val = '2021-04-01', '2023-04-01', '2021-04-01'
        st = _ratchet_storage(start, end, years=2, inventory_loss=0.0005, cmdty_consumed_inject=0.01)
        factors_args = dict(multi=([0.0, 2.5, 16.2, 45.0, 91.0], [0.14, 0.21, 0.64, 0.45, 0.85], 0.3))
        basis = full_polynomial_basis(5, 3); inv = 0.0
    elif name == 'c5':  # battery-style, two-factor, half-hourly one month
        N = num_sims or 2_000_000
        start, end, val = '2021-04-01 00:00', '2021-05-01 00:00', '2021-04-01 00:00'
        st = CmdtyStorage('30min', start, end, 0.0, 0.0, min_inventory=0.0, max_inventory=4.0, max_injection_rate=1.0,
                          max_withdrawal_rate=1.0, inventory_loss=0.002, cmdty_consumed_inject=0.05, cmdty_consumed_withdraw=0.05)
        factors_args = dict(multi=([0.0, 150.0], [0.25, 2.5], 0.4)); basis = '1 + x0 + x1 + x0**2 + x1**2 + x0*x1 + s + s**2 + s*x0 + s*x1'
        inv = 2.0
    else:
        raise ValueError("unknown config " + name)
    if start_inventory is not None:
        inv = start_inventory
    if basis_override is not None:
        basis = basis_override
    if storage_override is not None:
        st = storage_override
    elif end_override is not None:
        if name not in ('c3', 'c4'):
            raise ValueError("end override is only wired for the ratchet configs")
        st = _ratchet_storage(start, end_override, years=2 if name == 'c4' else 1, inventory_loss=0.0005, cmdty_consumed_inject=0.01)
    freq = st.freq
    pfreq = st.start.freq
    index = pd.period_range(st.start, st.end, freq=pfreq)
    val_period = pd.Period(val, freq=pfreq)
    fwd = _sin_forward(index, val_period)
    day_index = pd.period_range(pd.Period(val, freq='D'), pd.Period(val, freq='D') + 365 * 3, freq='D')
    rates = pd.Series(0.055, index=day_index)   # LSMCTest:148
    if 'seasonal' in factors_args:
        mr, sv, lv, ssv = factors_args['seasonal']
        factors, corrs = create_3_factor_season_params(freq, mr, sv, lv, ssv, val_period, st.end)
    elif 'one_factor' in factors_args:
        mr, sv = factors_args['one_factor']
        factors, corrs = [(mr, pd.Series(sv, index=index))], np.array([[1.0]])
    else:
        mrs, vols, rho = factors_args['multi']
        factors = [(m, pd.Series(v, index=index)) for m, v in zip(mrs, vols)]
        F = len(mrs)
        corrs = np.array([[rho ** abs(i - j) for j in range(F)] for i in range(F)])
    # Per-path inventories accumulate rounding of a few ulps per step (ulp(1400) = 2.3e-13, ulp(7000) = 9e-13), which
    # trips the bang-bang tolerance branch (SH.cs:124-128,148-152) at the Python wrapper's default 1e-12
    # (multi_factor.py:341); the reference's own error message recommends raising the tolerance.  The C# default is
    # 1e-10 (LVP.cs:86); the synthetic configs use 1e-9.
    tol = 1e-9
    return flatten_problem(st, val_period, inv, fwd, rates, _settle_20th_next_month, factors, corrs, N, basis, True,
                           numerical_tolerance=tol, **({} if grid_points is None else {"num_inventory_grid_points": grid_points}))


def full_polynomial_basis(num_factors, degree):
    """All monomials of total degree <= `degree` in x0..x{F-1} (C(F+degree, degree) terms; 56 for F=5, degree 3)."""
    terms = []

    def rec(f, remaining, powers):
        if f == num_factors:
            terms.append(list(powers))
            return
        for q in range(remaining + 1):
            rec(f + 1, remaining - q, powers + [q])
    rec(0, degree, [])
    terms.sort(key=lambda t: (sum(t), [-x for x in t]))
    out = []
    for t in terms:
        parts = ['x%d**%d' % (f, q) if q > 1 else 'x%d' % f for f, q in enumerate(t) if q > 0]
        out.append('*'.join(parts) if parts else '1')
    return ' + '.join(out)
