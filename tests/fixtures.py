"""Problem builders shared by the tests: the reference's own LSMC test fixture
(tests/Cmdty.Storage.Test/Lsmc/LsmcStorageValuationTest.cs:66-202) and its Black-76 call-like storage
(TestHelper.cs:86-216), flattened through the public host API."""
import math

import numpy as np
import pandas as pd

from storage_b200 import CmdtyStorage, RatchetInterp
from storage_b200._abi import Problem, CONSTRAINT_CONSTANT
from storage_b200.flatten import flatten_problem

STORAGE_START, STORAGE_END, VAL_DATE = "2019-08-03", "2020-04-01", "2019-08-29"
ONE_FACTOR_BASIS = "1 + x0 + x0**2"          # Ones + AllMarkovFactorAllPositiveIntegerPowersUpTo(2, 1)
TWO_FACTOR_BASIS = "1 + x0 + x0**2 + x1 + x1**2"


def settle_rule(period):   # LSMCTest:152: 20th of the following month
    return (period.asfreq("M") + 1).asfreq("D", "s") + 19


def simple_daily_storage(terminal=None):   # LSMCTest:77-89,129-142
    return CmdtyStorage("D", STORAGE_START, STORAGE_END, 1.25, 0.93, min_inventory=0.0, max_inventory=52_500.0,
                        max_injection_rate=625.0, max_withdrawal_rate=850.0, terminal_storage_npv=terminal)


def ratchet_daily_storage():   # LSMCTest:91-125
    ratchets = [
        (STORAGE_START, [(0.0, -702.7, 650.0), (15_000, -785.0, 552.5), (30_000, -790.6, 512.8), (40_000, -825.6, 498.6),
                         (52_500, -850.4, 480.0)]),
        ("2020-02-01", [(0.0, -645.35, 650.0), (13_000, -656.0, 552.5), (28_000, -689.6, 512.8), (42_000, -701.06, 498.6),
                        (52_500, -718.04, 480.0)]),
    ]
    return CmdtyStorage("D", STORAGE_START, STORAGE_END, 1.25, 0.93, ratchets=ratchets, ratchet_interp=RatchetInterp.LINEAR)


def sin_forward_curve(val_date=VAL_DATE, end=STORAGE_END):   # LSMCTest:154-160
    idx = pd.period_range(val_date, end, freq="D")
    days = np.arange(len(idx), dtype=np.float64)
    return pd.Series(53.5 + np.sin(2.0 * math.pi / 365.0 * days) * 24.6, index=idx)


def flat_rates(rate=0.055, start=VAL_DATE, days=600):
    return pd.Series(rate, index=pd.period_range(start, periods=days, freq="D"))


def reference_problem(storage=None, val_date=VAL_DATE, inventory=5_685.0, num_sims=2_000, mean_reversion=12.5,
                      spot_vol=0.95, basis=ONE_FACTOR_BASIS, discount_deltas=False, factors=None, corrs=None, **kw):
    """The `_1FactorParamsBuilder` fixture (LSMCTest:165-176)."""
    storage = storage or simple_daily_storage()
    curve_start = min(pd.Period(val_date, freq="D"), pd.Period(VAL_DATE, freq="D"))
    fwd = sin_forward_curve(curve_start)
    if factors is None:
        vols = pd.Series(spot_vol, index=pd.period_range(curve_start, STORAGE_END, freq="D"))
        factors, corrs = [(mean_reversion, vols)], np.array([[1.0]])
    return flatten_problem(storage, val_date, inventory, fwd, flat_rates(start=curve_start), settle_rule, factors, corrs, num_sims, basis,
                           discount_deltas, numerical_tolerance=kw.pop("numerical_tolerance", 1e-10), **kw)


def norm_cdf(x):
    return 0.5 * math.erfc(-x / math.sqrt(2.0))


def three_calls_problem(num_sims=2_000):
    """TestHelper.CreateThreeCallsLikeStorageTestData (TestHelper.cs:109-216) + LSMCTest:311-339, built directly as a
    flattened Problem: the withdrawal cost is `strike * volume` settled on the month's settlement date, which the C-ABI
    carries as per-period withdraw_cost / withdraw_cost_df.  Returns (problem, expected Black-76 value, option deltas)."""
    val = pd.Period("2019-08-29", freq="D")
    start, end = pd.Period("2019-12-01", freq="D"), pd.Period("2020-04-01", freq="D")
    curve_idx = pd.period_range(val, end, freq="D")
    i = np.arange(len(curve_idx), dtype=np.float64)
    fwd_curve = pd.Series(53.5 + np.sin(2.0 * math.pi / 365.0 * i) * 24.6, index=curve_idx)     # TestHelper.cs:86-107
    vol_curve = pd.Series(0.78 + np.sin(2.0 * math.pi / 365.0 * i) * 0.35, index=curve_idx)
    mean_reversion, rate = 16.5, 0.09
    dates = [pd.Period("2019-12-15", freq="D"), pd.Period("2020-01-20", freq="D"), pd.Period("2020-03-31", freq="D")]
    notionals = [1200.0, 800.0, 900.0]
    strikes = [fwd_curve[dates[0]], fwd_curve[dates[1]] + 2.0, fwd_curve[dates[2]] + 2.8]
    settle = {"2019-12": pd.Period("2020-01-20", freq="D"), "2020-01": pd.Period("2020-02-18", freq="D"),
              "2020-02": pd.Period("2020-03-21", freq="D"), "2020-03": pd.Period("2020-04-22", freq="D")}
    inventory = sum(notionals)
    periods = pd.period_range(start, end, freq="D")
    P = len(periods)

    def df(day):   # StorageHelper.CreateAct65ContCompDiscounter (SH.cs:262-273)
        return 1.0 if day <= val else math.exp(-(day.ordinal - val.ordinal) / 365.0 * rate)

    settle_df = np.ones(P)
    wd_cost = np.zeros(P)
    kinds, firsts, counts, lo, hi = [], [], [], [], []
    cid = np.zeros(P, dtype=np.int32)
    rates = {0.0: 0}
    kinds.append(CONSTRAINT_CONSTANT); firsts.append(0); counts.append(1); lo.append(0.0); hi.append(0.0)
    for p, per in enumerate(periods[:-1]):
        settle_df[p] = df(settle[str(per.asfreq("M"))])
        if per in dates:
            k = dates.index(per)
            wd_cost[p] = strikes[k]
            kinds.append(CONSTRAINT_CONSTANT); firsts.append(len(lo)); counts.append(1); lo.append(-notionals[k]); hi.append(0.0)
            cid[p] = len(kinds) - 1
    cid[P - 1] = cid[P - 2]
    sim_time = np.array([(per.ordinal - val.ordinal) / 365.0 for per in periods])
    zeros = np.zeros(P)
    pb = Problem(
        n_periods=P, n_factors=1, n_basis=3, n_sims=num_sims, extra_decisions=0, first_period_is_current=0,
        discount_deltas=0, must_be_empty_at_end=0, numerical_tolerance=1e-10, start_inventory=inventory,
        grid_spacing=inventory / 99.0, terminal_coef=(0.0, 0.0, 0.0, 0.0),
        sim_time=sim_time, mean_reversion=np.array([mean_reversion]), factor_vol=vol_curve[periods].values.reshape(P, 1),
        factor_corr=np.array([[1.0]]), forward=fwd_curve[periods].values,
        spot_power=np.zeros(3, dtype=np.int32), factor_power=np.array([[0], [1], [2]], dtype=np.int32),
        min_inventory=zeros.copy(), max_inventory=np.full(P, inventory), constraint_id=cid,
        inject_cost=zeros.copy(), inject_cost_df=np.ones(P), withdraw_cost=wd_cost, withdraw_cost_df=settle_df.copy(),
        consumed_inject=zeros.copy(), consumed_withdraw=zeros.copy(), inventory_loss=zeros.copy(),
        inventory_cost=zeros.copy(), inventory_cost_df=np.ones(P), settle_df=settle_df,
        constraint_kind=np.array(kinds, dtype=np.int32), constraint_first=np.array(firsts, dtype=np.int32),
        constraint_count=np.array(counts, dtype=np.int32),
        pillar_inventory=np.zeros(len(lo)), pillar_min_rate=np.array(lo), pillar_max_rate=np.array(hi),
        periods=periods,
    ).finalize()

    # Black-76 (TestHelper.cs:38-86)
    expected, deltas = 0.0, {}
    for d, k, n in zip(dates, strikes, notionals):
        t = (d.ordinal - val.ordinal) / 365.0
        implied = vol_curve[d] * math.sqrt((1 - math.exp(-2 * mean_reversion * t)) / 2.0 / mean_reversion / t)
        f = fwd_curve[d]
        vrt = implied * math.sqrt(t)
        d1 = (math.log(f / k) + implied * implied / 2 * t) / vrt
        expected += df(settle[str(d.asfreq("M"))]) * (f * norm_cdf(d1) - k * norm_cdf(d1 - vrt)) * n
        deltas[int(np.where(periods == d)[0][0])] = norm_cdf(d1) * n
    return pb, expected, deltas


def forced_inject_withdraw_problem(num_sims=1_000):
    """Calculate_StorageWithForcedInjectAndWithdraw_NpvAlmostEqualsTrivialIntrinsicCalc (LSMCTest:686-837), built directly as
    a flattened Problem.  The reference builds the storage with ...RatesPolynomial, but every one of its five tables has
    two rows with equal rates, which AddInjectWithdrawRanges collapses to ConstantInjectWithdrawConstraint
    (CmdtyStorageBuilderExtensions.cs:163-171): no polynomial is ever fitted.  Injection costs settle 10 days after the
    injection, withdrawal costs 4 days after (per-period cost discount factors of the C-ABI).
    Returns (problem, expected NPV of the forced schedule valued on the forward curve)."""
    val = pd.Period("2019-08-29", freq="D")
    start, end = pd.Period("2019-12-01", freq="D"), pd.Period("2020-04-01", freq="D")
    max_inventory = 10_000.0
    inj_rate, inj_days, inj_start = 211.5, 20, pd.Period("2019-12-20", freq="D")
    wd_rate, wd_days, wd_start = 187.54, 15, pd.Period("2020-02-05", freq="D")
    inj_cost, inj_consumed, wd_cost, wd_consumed = 1.23, 0.01, 0.98, 0.015
    mean_reversion, rate = 16.5, 0.09
    curve_idx = pd.period_range(val, end, freq="D")
    i = np.arange(len(curve_idx), dtype=np.float64)
    fwd_curve = pd.Series(53.5 + np.sin(2.0 * math.pi / 365.0 * i) * 24.6, index=curve_idx)     # TestHelper.cs:86-107
    vol_curve = pd.Series(0.78 + np.sin(2.0 * math.pi / 365.0 * i) * 0.35, index=curve_idx)
    settle = {"2019-12": pd.Period("2020-01-20", freq="D"), "2020-01": pd.Period("2020-02-18", freq="D"),
              "2020-02": pd.Period("2020-03-21", freq="D"), "2020-03": pd.Period("2020-04-22", freq="D")}
    periods = pd.period_range(start, end, freq="D")
    P = len(periods)

    def df(day):
        return 1.0 if day <= val else math.exp(-(day.ordinal - val.ordinal) / 365.0 * rate)

    def forced(per):
        if inj_start <= per < inj_start + inj_days:
            return inj_rate
        if wd_start <= per < wd_start + wd_days:
            return -wd_rate
        return 0.0

    rates = [0.0, inj_rate, -wd_rate]
    cid = np.array([rates.index(forced(per)) for per in periods], dtype=np.int32)
    settle_df, inj_df, wd_df = np.ones(P), np.ones(P), np.ones(P)
    for p, per in enumerate(periods[:-1]):
        settle_df[p] = df(settle[str(per.asfreq("M"))])
        inj_df[p], wd_df[p] = df(per + 10), df(per + 4)
    zeros = np.zeros(P)
    K_spot, K_fac = np.zeros(3, dtype=np.int32), np.array([[0], [1], [2]], dtype=np.int32)
    pb = Problem(
        n_periods=P, n_factors=1, n_basis=3, n_sims=num_sims, extra_decisions=0, first_period_is_current=0,
        discount_deltas=0, must_be_empty_at_end=0, numerical_tolerance=1e-10, start_inventory=0.0,
        grid_spacing=max_inventory / 499.0, terminal_coef=(0.0, 0.0, 0.0, 0.0),
        sim_time=np.array([(per.ordinal - val.ordinal) / 365.0 for per in periods]), mean_reversion=np.array([mean_reversion]),
        factor_vol=vol_curve[periods].values.reshape(P, 1), factor_corr=np.array([[1.0]]), forward=fwd_curve[periods].values,
        spot_power=K_spot, factor_power=K_fac,
        min_inventory=zeros.copy(), max_inventory=np.full(P, max_inventory), constraint_id=cid,
        inject_cost=np.full(P, inj_cost), inject_cost_df=inj_df, withdraw_cost=np.full(P, wd_cost), withdraw_cost_df=wd_df,
        consumed_inject=np.full(P, inj_consumed), consumed_withdraw=np.full(P, wd_consumed), inventory_loss=zeros.copy(),
        inventory_cost=zeros.copy(), inventory_cost_df=np.ones(P), settle_df=settle_df,
        constraint_kind=np.full(3, CONSTRAINT_CONSTANT, dtype=np.int32), constraint_first=np.arange(3, dtype=np.int32),
        constraint_count=np.ones(3, dtype=np.int32), pillar_inventory=np.zeros(3), pillar_min_rate=np.array(rates),
        pillar_max_rate=np.array(rates), periods=periods,
    ).finalize()
    expected = 0.0
    for k in range(inj_days):
        day = inj_start + k
        expected += -fwd_curve[day] * inj_rate * (1 + inj_consumed) * df(settle[str(day.asfreq("M"))]) - inj_cost * inj_rate * df(day + 10)
    for k in range(wd_days):
        day = wd_start + k
        expected += fwd_curve[day] * wd_rate * (1 - wd_consumed) * df(settle[str(day.asfreq("M"))]) - wd_cost * wd_rate * df(day + 4)
    return pb, expected
