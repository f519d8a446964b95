"""Flatten (storage, curves, model) into the C-ABI `lsmc_problem` tables.

This is the part of `_net_multi_factor_calc` (multi_factor.py:386-433) and of the head of
`LsmcStorageValuation.Calculate` (LSV.cs:61-95,131-143,222-223) that deals in dates and delegates; the
C# shim described in INTEGRATION.md does exactly the same with the real .NET objects.
"""
import math
import typing as tp

import numpy as np
import pandas as pd

from . import utils
from ._abi import Problem, CONSTRAINT_CONSTANT
from .cmdty_storage import CmdtyStorage


class TrivialResult(tp.NamedTuple):
    """Expired / at-end-period exits that never reach the native library (LSV.cs:64-84)."""
    npv: float
    expired: bool


def act65_cont_comp_discount_factors(interest_rates: pd.Series, present_day: pd.Period,
                                     cash_flow_days: tp.Sequence[pd.Period]) -> np.ndarray:
    """StorageHelper.CreateAct65ContCompDiscounterFromSeries (SH.cs:251-273)."""
    if not isinstance(interest_rates.index, pd.PeriodIndex):
        interest_rates = pd.Series(interest_rates.values, index=pd.PeriodIndex(interest_rates.index, freq="D"))
    out = np.ones(len(cash_flow_days))
    ords = np.array([d.ordinal for d in cash_flow_days], dtype=np.int64)
    later = ords > present_day.ordinal
    if later.any():
        days = pd.PeriodIndex.from_ordinals(ords[later], freq="D")
        idx = interest_rates.index.get_indexer(days)
        if (idx < 0).any():
            raise ValueError("No interest rate provided for {}.".format(days[int(np.argmax(idx < 0))]))
        rates = interest_rates.values[idx].astype(np.float64)
        offs = (ords[later] - present_day.ordinal).astype(np.float64)
        out[later] = np.array([math.exp(-o / 365.0 * r) for o, r in zip(offs, rates)])
    return out


def flatten_problem(cmdty_storage: CmdtyStorage,
                    val_date,
                    inventory: float,
                    fwd_curve: pd.Series,
                    interest_rates: pd.Series,
                    settlement_rule: tp.Callable[[pd.Period], tp.Any],
                    factors: tp.Sequence[tp.Tuple[float, utils.CurveType]],
                    factor_corrs: np.ndarray,
                    num_sims: int,
                    basis_funcs: str,
                    discount_deltas: bool,
                    extra_decisions: tp.Optional[int] = None,
                    num_inventory_grid_points: int = 100,
                    numerical_tolerance: float = 1e-12) -> tp.Union[Problem, TrivialResult]:
    freq = cmdty_storage.freq
    current = utils.to_period(val_date, freq)
    if inventory < 0:    # LSV.cs:61
        raise ValueError("Inventory cannot be negative.")
    if current > cmdty_storage.end:    # LSV.cs:64-68
        return TrivialResult(0.0, True)
    if current == cmdty_storage.end:    # LSV.cs:70-84
        if cmdty_storage.empty_at_end:
            if inventory > 0:
                raise ValueError("Storage must be empty at end, but inventory is greater than zero.")
            return TrivialResult(0.0, True)
        spot = utils.curve_lookup(fwd_curve, current, freq, "forward price")
        return TrivialResult(cmdty_storage.terminal_storage_npv(spot, inventory), False)

    start_active = max(cmdty_storage.start, current)    # SH.cs:45
    periods = pd.period_range(start=start_active, end=cmdty_storage.end, freq=start_active.freq)
    P = len(periods)
    first_is_current = int(start_active == current)

    # forward curve coverage (LSV.cs:91-95)
    if not isinstance(fwd_curve.index, pd.PeriodIndex):
        raise ValueError("fwd_curve must be indexed by a pandas PeriodIndex")
    if fwd_curve.index[0] > start_active:
        raise ValueError("Forward curve starts too late. Must start on or before the period {}.".format(start_active))
    if fwd_curve.index[-1] < cmdty_storage.end:
        raise ValueError("Forward curve does not extend until storage end period.")
    forward = utils.series_values_for(fwd_curve, periods, "forward price")

    F = len(factors)
    K_spot, K_fac = utils.parse_basis_functions(basis_funcs, F)
    factor_vol = np.zeros((P, F))
    mean_reversion = np.zeros(F)
    for f, (mr, vol_curve) in enumerate(factors):
        mean_reversion[f] = float(mr)
        lo = first_is_current    # vols are only needed for simulated delivery periods
        if isinstance(vol_curve, pd.Series):
            factor_vol[lo:, f] = utils.series_values_for(vol_curve, periods[lo:], "factor {} vol".format(f))
        else:
            for p in range(lo, P):
                factor_vol[p, f] = utils.curve_lookup(vol_curve, periods[p], freq, "factor {} vol".format(f))

    current_start = current.start_time
    sim_time = np.array([utils.act_365_datetime(current_start, per.start_time) for per in periods])

    present_day = utils.first_day(current)    # LSV.cs:131
    settle_days = [utils.to_day(settlement_rule(per)) for per in periods[:-1]]
    settle_df = np.ones(P)
    settle_df[:-1] = act65_cont_comp_discount_factors(interest_rates, present_day, settle_days)
    cost_days = [utils.first_day(per) for per in periods[:-1]]    # CmdtyStorage.cs:337-338: period.First<Day>()
    cost_df = np.ones(P)
    cost_df[:-1] = act65_cont_comp_discount_factors(interest_rates, present_day, cost_days)

    def per_period(fn, last=0.0):
        out = np.full(P, last, dtype=np.float64)
        for p in range(P - 1):
            out[p] = fn(periods[p])
        return out

    min_inv = np.array([cmdty_storage.min_inventory(per) for per in periods])
    max_inv = np.array([cmdty_storage.max_inventory(per) for per in periods])

    # constraint tables, de-duplicated
    table: tp.Dict[tp.Any, int] = {}
    kinds, firsts, counts, p_inv, p_min, p_max = [], [], [], [], [], []
    constraint_id = np.zeros(P, dtype=np.int32)
    for p in range(P - 1):
        c = cmdty_storage.constraint_for(periods[p])
        key = (c.kind, c.inventory, c.min_rate, c.max_rate) if c.kind != CONSTRAINT_CONSTANT else (c.kind, c.min_rate[0], c.max_rate[0])
        if key not in table:
            table[key] = len(kinds)
            kinds.append(c.kind)
            firsts.append(len(p_inv))
            if c.kind == CONSTRAINT_CONSTANT:
                counts.append(1)
                p_inv.append(0.0); p_min.append(c.min_rate[0]); p_max.append(c.max_rate[0])
            else:
                counts.append(len(c.inventory))
                p_inv.extend(c.inventory); p_min.extend(c.min_rate); p_max.extend(c.max_rate)
        constraint_id[p] = table[key]
    constraint_id[P - 1] = constraint_id[P - 2]

    # FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange (:65-82)
    if num_inventory_grid_points < 3:
        raise ValueError("Parameter numGridPointsOverGlobalInventoryRange value must be at least 3.")
    storage_periods = pd.period_range(start=cmdty_storage.start, end=cmdty_storage.end, freq=start_active.freq)
    global_max = max(cmdty_storage.max_inventory(per) for per in _distinct_constraint_periods(cmdty_storage, storage_periods))
    global_min = min(cmdty_storage.min_inventory(per) for per in _distinct_constraint_periods(cmdty_storage, storage_periods))
    grid_spacing = (global_max - global_min) / (num_inventory_grid_points - 1)

    return Problem(
        n_periods=P, n_factors=F, n_basis=len(K_spot), n_sims=int(num_sims),
        extra_decisions=int(extra_decisions or 0), first_period_is_current=first_is_current,
        discount_deltas=int(bool(discount_deltas)), must_be_empty_at_end=int(cmdty_storage.empty_at_end),
        numerical_tolerance=float(numerical_tolerance), start_inventory=float(inventory), grid_spacing=float(grid_spacing),
        terminal_coef=cmdty_storage.terminal_coefficients(),
        sim_time=sim_time, mean_reversion=mean_reversion, factor_vol=factor_vol,
        factor_corr=np.asarray(factor_corrs, dtype=np.float64), forward=forward,
        spot_power=K_spot, factor_power=K_fac,
        min_inventory=min_inv, max_inventory=max_inv, constraint_id=constraint_id,
        inject_cost=per_period(cmdty_storage.unit_injection_cost), inject_cost_df=cost_df,
        withdraw_cost=per_period(cmdty_storage.unit_withdrawal_cost), withdraw_cost_df=cost_df.copy(),
        consumed_inject=per_period(cmdty_storage.pcnt_consumed_inject),
        consumed_withdraw=per_period(cmdty_storage.pcnt_consumed_withdraw),
        inventory_loss=per_period(cmdty_storage.inventory_pcnt_loss),
        inventory_cost=per_period(cmdty_storage.unit_inventory_cost), inventory_cost_df=cost_df.copy(),
        settle_df=settle_df,
        constraint_kind=np.array(kinds, dtype=np.int32), constraint_first=np.array(firsts, dtype=np.int32),
        constraint_count=np.array(counts, dtype=np.int32),
        pillar_inventory=np.array(p_inv), pillar_min_rate=np.array(p_min), pillar_max_rate=np.array(p_max),
        periods=periods,
    ).finalize()


def _distinct_constraint_periods(storage: CmdtyStorage, periods: pd.PeriodIndex):
    """All storage periods (the reference takes the max/min over every one, FixedSpacing…cs:75-79)."""
    return periods
