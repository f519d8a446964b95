"""multi_factor_value / three_factor_seasonal_value — the reference's Python entry points for the LSMC path
(src/Cmdty.Storage.Python/cmdty_storage/multi_factor.py:324-455) with the same signatures and result tuple, running
the valuation on a B200 through the C ABI (include/lsmc_b200.h) instead of pythonnet -> LsmcStorageValuation.Calculate.

What differs from the reference, by necessity:
  * random numbers: seeded calls draw the reference's own stream (MT19937 + MathNet Normal.Samples + antithetic, as
    MersenneTwisterGenerator(seed, true) feeds Cmdty.Core's simulator, LVP.cs:185-189; csrc/mt19937.h) on the host and
    reproduce the reference's seeded golden NPVs; unseeded calls (and SEEDED_RNG = "philox") use the device
    Philox4x32-10 stream, which agrees with the reference to Monte-Carlo error only;
  * `intrinsic_npv` / `intrinsic_profile` come from the library's host routine lsmc_intrinsic (storage_b200/intrinsic.py),
    the reference's IntrinsicStorageValuation restated (pinned by the reference's golden value 1703773.0757192627);
  * there is no CPU fallback: without liblsmc_b200.so or a B200 the call raises.
"""
import logging
import os
import typing as tp
from datetime import date

import numpy as np
import pandas as pd

from . import _abi, _native, utils
from .cmdty_storage import CmdtyStorage
from .flatten import flatten_problem, TrivialResult
from .intrinsic import intrinsic_of_problem
from .multi_factor_params import create_3_factor_season_params

logger: logging.Logger = logging.getLogger('cmdty.storage.multi-factor')

FactorCorrsType = tp.Optional[tp.Union[float, np.ndarray]]


class TriggerPricePoint(tp.NamedTuple):
    volume: float
    price: float


class TriggerPriceProfile(tp.NamedTuple):
    inject_triggers: tp.List[TriggerPricePoint]
    withdraw_triggers: tp.List[TriggerPricePoint]


class MultiFactorValuationResults(tp.NamedTuple):
    npv: float
    deltas: pd.Series
    expected_profile: pd.DataFrame
    intrinsic_npv: float
    intrinsic_profile: tp.Optional[pd.DataFrame]
    sim_spot_regress: pd.DataFrame
    sim_spot_valuation: pd.DataFrame
    sim_inventory: pd.DataFrame
    sim_inject_withdraw: pd.DataFrame
    sim_cmdty_consumed: pd.DataFrame
    sim_inventory_loss: pd.DataFrame
    sim_net_volume: pd.DataFrame
    sim_pv: pd.DataFrame
    trigger_prices: pd.DataFrame
    trigger_profiles: pd.Series

    @property
    def extrinsic_npv(self):
        return self.npv - self.intrinsic_npv


def _validate_multi_factor_params(factors, factor_corrs: FactorCorrsType) -> np.ndarray:
    """The checks of the reference's validator (multi_factor.py:112-146) in the same order and with the same messages — the
    error strings are what its wrapper tests assert — on the whole matrix at once; the first offender in row-major order is
    the one reported, as in the reference's element loop."""
    n = len(factors)
    if n == 0:
        raise ValueError("factors cannot be empty.")
    if n == 1 and factor_corrs is None:
        corrs = np.ones((1, 1))
    elif n == 2 and isinstance(factor_corrs, (float, int)):
        corrs = np.array([[1.0, float(factor_corrs)], [float(factor_corrs), 1.0]])
    else:
        corrs = np.asarray(factor_corrs)
    if corrs.ndim != 2:
        raise ValueError("Factor correlation matrix is not 2-dimensional.")
    if corrs.shape[0] != corrs.shape[1]:
        raise ValueError("Factor correlation matrix is not square.")
    corrs = corrs.astype(np.float64)
    diagonal = np.eye(corrs.shape[0], dtype=bool)
    offending = np.where(diagonal, ~np.isclose(corrs, 1.0), ~((corrs >= -1.0) & (corrs <= 1.0)))
    if offending.any():
        i, j = (int(v) for v in np.argwhere(offending)[0])
        if i == j:
            raise ValueError("Factor correlation on diagonal position ({i}, {j}) value of {corr} not valid as not "
                             "equal to 1.".format(i=i, j=j, corr=corrs[i, j]))
        raise ValueError("Factor correlation in position ({i}, {j}) value of {corr} not valid as not in the "
                         "interval [-1, 1]".format(i=i, j=j, corr=corrs[i, j]))
    if n != corrs.shape[0]:
        raise ValueError("factor_corrs and factors must be of the same length. factor_corrs is of length {} and "
                         "factors is of length {}.".format(corrs.shape[0], n))
    return corrs


def _seeds(seed, fwd_sim_seed):
    """MersenneTwisterGenerator(seed, antithetic: true) twice, or one generator continued (LVP.cs:181-191) -> two
    Philox keys; without fwd_sim_seed the valuation set uses a fixed distinct sub-stream of the same seed."""
    if seed is None:
        seed = int.from_bytes(os.urandom(8), "little")
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    if fwd_sim_seed is None:
        fwd = (seed + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    else:
        fwd = int(fwd_sim_seed) & 0xFFFFFFFFFFFFFFFF
    return seed, fwd


# Which draw stream a SEEDED call uses.  "mt19937": the reference's own (MathNet MT19937 + Normal.Samples, generated serially
# on the host, csrc/mt19937.h): the same seed gives the reference's numbers (its golden NPVs 1780380.7581833513 and
# 1766460.137569665 to about 1e-12, test_multi_factor.py:226,298).  "philox": the device stream, for seeded runs with
# millions of paths where a serial host generator (about 15 ns per normal) would dominate.  Unseeded calls always use Philox.
SEEDED_RNG = "mt19937"


def _rngs(seed, fwd_sim_seed):
    """multi_factor.py:436-440 / LVP.cs:181-191: regression generator from `seed`; the valuation generator is a new one from
    `fwd_sim_seed`, or the regression generator continuing its sequence when that is None."""
    if seed is None or SEEDED_RNG != "mt19937":
        reg_seed, val_seed = _seeds(seed, fwd_sim_seed)
        return _abi.Rng.philox(reg_seed, antithetic=True), _abi.Rng.philox(val_seed, antithetic=True)
    reg = _abi.Rng.mt19937(seed, antithetic=True)
    val = _abi.Rng.mt19937_continue(antithetic=True) if fwd_sim_seed is None else _abi.Rng.mt19937(fwd_sim_seed, antithetic=True)
    return reg, val


def three_factor_seasonal_value(cmdty_storage: CmdtyStorage,
                                val_date: utils.TimePeriodSpecType,
                                inventory: float,
                                fwd_curve: pd.Series,
                                interest_rates: pd.Series,
                                settlement_rule: tp.Callable[[pd.Period], date],
                                spot_mean_reversion: float,
                                spot_vol: float,
                                long_term_vol: float,
                                seasonal_vol: float,
                                num_sims: int,
                                basis_funcs: str,
                                discount_deltas: bool,
                                seed: tp.Optional[int] = None,
                                fwd_sim_seed: tp.Optional[int] = None,
                                extra_decisions: tp.Optional[int] = None,
                                num_inventory_grid_points: int = 100,
                                numerical_tolerance: float = 1E-12,
                                on_progress_update: tp.Optional[tp.Callable[[float], None]] = None,
                                ) -> MultiFactorValuationResults:
    current = utils.to_period(val_date, cmdty_storage.freq)
    factors, factor_corrs = create_3_factor_season_params(cmdty_storage.freq, spot_mean_reversion, spot_vol, long_term_vol,
                                                          seasonal_vol, current, cmdty_storage.end)
    # Transform factors x_st -> x0, x_lt -> x1, x_sw -> x2 (multi_factor.py:349-350)
    basis_func_transformed = basis_funcs.replace('x_st', 'x0').replace('x_lt', 'x1').replace('x_sw', 'x2')
    return _lsmc_calc(cmdty_storage, fwd_curve, interest_rates, inventory, factors, factor_corrs, num_inventory_grid_points,
                      num_sims, numerical_tolerance, on_progress_update, basis_func_transformed, seed, fwd_sim_seed,
                      settlement_rule, val_date, discount_deltas, extra_decisions)


def multi_factor_value(cmdty_storage: CmdtyStorage,
                       val_date: utils.TimePeriodSpecType,
                       inventory: float,
                       fwd_curve: pd.Series,
                       interest_rates: pd.Series,
                       settlement_rule: tp.Callable[[pd.Period], date],
                       factors: tp.Iterable[tp.Tuple[float, utils.CurveType]],
                       factor_corrs: FactorCorrsType,
                       num_sims: int,
                       basis_funcs: str,
                       discount_deltas: bool,
                       seed: tp.Optional[int] = None,
                       fwd_sim_seed: tp.Optional[int] = None,
                       extra_decisions: tp.Optional[int] = None,
                       num_inventory_grid_points: int = 100,
                       numerical_tolerance: float = 1E-12,
                       on_progress_update: tp.Optional[tp.Callable[[float], None]] = None,
                       ) -> MultiFactorValuationResults:
    factors = list(factors)
    factor_corrs = _validate_multi_factor_params(factors, factor_corrs)
    if utils.pandas_freq(cmdty_storage.freq) != utils.pandas_freq_of_index(fwd_curve.index):
        raise ValueError("cmdty_storage and forward_curve have different frequencies.")
    return _lsmc_calc(cmdty_storage, fwd_curve, interest_rates, inventory, factors, factor_corrs, num_inventory_grid_points,
                      num_sims, numerical_tolerance, on_progress_update, basis_funcs, seed, fwd_sim_seed, settlement_rule,
                      val_date, discount_deltas, extra_decisions)


# Overridable for tests / multi-GPU hosts: CUDA device index used by the wrappers.
DEVICE = 0
# The reference always returns the eight sims x periods panels; at millions of paths they dominate the wall time
# (device -> host -> DataFrame).  Set to False to get empty frames instead.
RETURN_SIM_PANELS = True


def _lsmc_calc(cmdty_storage, fwd_curve, interest_rates, inventory, factors, factor_corrs, num_inventory_grid_points, num_sims,
               numerical_tolerance, on_progress_update, basis_funcs, seed, fwd_sim_seed, settlement_rule, val_date,
               discount_deltas, extra_decisions) -> MultiFactorValuationResults:
    """_net_multi_factor_calc (multi_factor.py:386-455) without .NET: flatten -> lsmc_calculate -> pandas."""
    problem = flatten_problem(cmdty_storage, val_date, inventory, fwd_curve, interest_rates, settlement_rule, factors,
                              factor_corrs, num_sims, basis_funcs, discount_deltas, extra_decisions, num_inventory_grid_points,
                              numerical_tolerance)
    freq = utils.pandas_freq(cmdty_storage.freq)
    if isinstance(problem, TrivialResult):    # LSV.cs:64-84: expired / at the end period, never reaches the device
        empty_index = pd.PeriodIndex([], freq=freq)
        empty = pd.DataFrame(index=empty_index)
        profile = pd.DataFrame(index=empty_index, columns=_PROFILE_COLUMNS, dtype=np.float64)
        if on_progress_update is not None:
            on_progress_update(1.0)
        intrinsic = intrinsic_of_problem(problem)
        return MultiFactorValuationResults(problem.npv, pd.Series(index=empty_index, dtype=np.float64), profile, intrinsic.npv,
                                           intrinsic.profile,
                                           empty, empty, empty, empty, empty, empty, empty, empty,
                                           pd.DataFrame(index=empty_index, columns=_TRIGGER_COLUMNS, dtype=np.float64),
                                           pd.Series(index=empty_index, dtype=object))
    periods = problem.periods
    P, S, N = problem.n_periods, problem.n_sim_periods, int(problem.n_sims)
    logger.info('Calculating intrinsic value.')
    intrinsic = intrinsic_of_problem(problem)
    logger.info('Calculation of intrinsic value complete.')
    reg_rng, val_rng = _rngs(seed, fwd_sim_seed)
    panels = "all" if RETURN_SIM_PANELS else ()
    logger.info('Calculating LSMC value.')
    with _native.Context(DEVICE) as ctx:
        res = ctx.calculate(problem, reg_rng, val_rng, panels=panels, on_progress=on_progress_update)
        timings = ctx.timings()
    logger.info('Calculation of LSMC value complete.')
    logger.info('device ms: regression sim %.2f, backward %.2f, valuation sim %.2f, forward %.2f',
                timings['regression_sim_ms'], timings['backward_ms'], timings['valuation_sim_ms'], timings['forward_ms'])

    deltas = pd.Series(res.deltas, index=periods)
    expected_profile = pd.DataFrame(res.profile, index=periods, columns=_PROFILE_COLUMNS)
    decision_periods = periods[:-1]
    ts = res.trigger_summary
    trigger_prices = pd.DataFrame({'inject_volume': ts[:, 0], 'inject_trigger_price': ts[:, 1],
                                   'withdraw_volume': ts[:, 2], 'withdraw_trigger_price': ts[:, 3]}, index=decision_periods)
    profiles = []
    for p in range(P - 1):
        inj = [TriggerPricePoint(v, pr) for v, pr in res.trigger_inject[p] if not np.isnan(v)]
        wdr = [TriggerPricePoint(v, pr) for v, pr in res.trigger_withdraw[p] if not np.isnan(v)]
        profiles.append(TriggerPriceProfile(inj, wdr))
    trigger_profiles = pd.Series(profiles, index=decision_periods, dtype=object)
    sim_periods = periods[P - S:]

    def frame(a, index):
        return pd.DataFrame(a, index=index) if a is not None else pd.DataFrame(index=index)

    return MultiFactorValuationResults(res.npv, deltas, expected_profile, intrinsic.npv, intrinsic.profile,
                                       frame(res.spot_regress, sim_periods), frame(res.spot_valuation, sim_periods),
                                       frame(res.inventory, periods), frame(res.inject_withdraw, periods),
                                       frame(res.consumed, periods), frame(res.loss, periods), frame(res.net_volume, periods),
                                       frame(res.pv, periods), trigger_prices, trigger_profiles)


_PROFILE_COLUMNS = ['inventory', 'inject_withdraw_volume', 'cmdty_consumed', 'inventory_loss', 'net_volume', 'period_pv']
_TRIGGER_COLUMNS = ['inject_volume', 'inject_trigger_price', 'withdraw_volume', 'withdraw_trigger_price']


class MultiFactorSpotSim:
    """multi_factor.py:49-92: spot simulation on its own, served by the K1 device kernel."""

    def __init__(self, freq: str, factors, factor_corrs: FactorCorrsType, current_date, fwd_curve: utils.CurveType,
                 sim_periods, seed: tp.Optional[int] = None, antithetic: bool = False):
        self._factors = list(factors)
        self._corrs = _validate_multi_factor_params(self._factors, factor_corrs)
        self._pfreq = utils.pandas_freq(freq)
        self._freq = freq
        self._periods = pd.PeriodIndex([utils.to_period(p, freq) for p in sim_periods], freq=self._pfreq)
        if len(self._periods) == 0:
            raise ValueError("sim_periods cannot be empty.")
        self._current = pd.Timestamp(current_date.start_time if isinstance(current_date, pd.Period) else current_date)
        self._fwd = fwd_curve
        self._given_seed = seed
        self._seed, _ = _seeds(seed, None)
        self._antithetic = bool(antithetic)
        self._calls = 0

    def simulate(self, num_sims: int) -> pd.DataFrame:
        S, F = len(self._periods), len(self._factors)
        sim_time = np.array([utils.act_365_datetime(self._current, p.start_time) for p in self._periods])
        if np.any(np.diff(np.concatenate([[0.0], sim_time])) <= 0):
            raise ValueError("sim_periods must be strictly increasing and after current_date.")
        forward = np.array([utils.curve_lookup(self._fwd, p, self._freq, "forward price") for p in self._periods])
        vol = np.array([[utils.curve_lookup(vc, p, self._freq, "factor vol") for _, vc in self._factors] for p in self._periods])
        z, o = np.zeros(S + 1), np.ones(S + 1)
        pad = lambda a: np.concatenate([a, a[-1:]])   # noqa: E731  a terminal dummy period so that P = S + 1 >= 2
        pb = _abi.Problem(
            n_periods=S + 1, n_factors=F, n_basis=1, n_sims=int(num_sims), first_period_is_current=0, must_be_empty_at_end=1,
            start_inventory=0.0, grid_spacing=1.0, sim_time=pad(sim_time) + np.concatenate([np.zeros(S), [1.0 / 365.0]]),
            mean_reversion=[float(m) for m, _ in self._factors], factor_vol=np.vstack([vol, vol[-1:]]), factor_corr=self._corrs,
            forward=pad(forward), spot_power=[0], factor_power=np.zeros((1, F), dtype=np.int32),
            min_inventory=z, max_inventory=z, constraint_id=np.zeros(S + 1, dtype=np.int32), inject_cost=z, inject_cost_df=o,
            withdraw_cost=z, withdraw_cost_df=o, consumed_inject=z, consumed_withdraw=z, inventory_loss=z, inventory_cost=z,
            inventory_cost_df=o, settle_df=o, constraint_kind=[_abi.CONSTRAINT_CONSTANT], constraint_first=[0],
            constraint_count=[1], pillar_inventory=[0.0], pillar_min_rate=[0.0], pillar_max_rate=[0.0]).finalize()
        with _native.Context(DEVICE) as ctx:
            ctx.set_problem(pb)
            # successive calls continue the stream like one MersenneTwisterGenerator instance would
            if self._given_seed is not None and SEEDED_RNG == "mt19937":
                # the reference's stream; successive calls continue it like one MersenneTwisterGenerator instance
                if self._calls == 0:
                    self._sims_drawn = 0
                # (the dummy terminal period of the problem below is not part of the reference's draw: S steps, padded with zeros)
                z = _native.mt19937_normals(self._given_seed, self._antithetic, self._sims_drawn + int(num_sims), S, F)[:, :, self._sims_drawn:]
                self._sims_drawn += int(num_sims)
                ctx.simulate(0, _abi.Rng.from_normals(np.ascontiguousarray(np.concatenate([z, np.zeros((1, F, int(num_sims)))]))))
            else:
                ctx.simulate(0, _abi.Rng.philox((self._seed + self._calls) & 0xFFFFFFFFFFFFFFFF, antithetic=self._antithetic))
            spot = ctx.sim_panel(0)
        self._calls += 1
        return pd.DataFrame(data=spot[:S], index=self._periods)
