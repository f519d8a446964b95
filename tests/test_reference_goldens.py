"""The reference's own SEEDED golden values, reproduced (src/Cmdty.Storage.Python/tests/test_multi_factor.py).

They are functions of MathNet's MT19937 stream and of Cmdty.Core's draw order, neither of which is in /root/reference; the draw
order was found by search against the twelve golden spot prices (scripts/mt19937_search.py) and then CONFIRMED by the two golden
NPVs, which the CPU oracle — simulator model, inventory space, grids, Householder-QR regression, decision sweep, forward pass —
reproduces to the last bit.  That pins the oracle end to end against numbers the reference itself asserts.
Host arithmetic only: runs without a GPU (the device path is held to the same numbers in tests/test_gpu_wrappers.py)."""
import numpy as np
import pandas as pd
import pytest

from storage_b200 import CmdtyStorage, _abi, _native, utils
from storage_b200._abi import Rng
from storage_b200.flatten import flatten_problem
from storage_b200.multi_factor_params import create_3_factor_season_params
from oracle import oracle
import mt19937_ref

GOLDEN_SPOT = np.array([[52.59976397688973, 46.1206448628463, 58.15838580682589, 65.500441945042979],       # test_multi_factor.py:74-91
                        [57.559631642935514, 72.0381089486175, 82.49607173562342, 42.812676607997183],      # [date][sim]
                        [89.40526992772634, 85.18869803117379, 138.68587285875978, 76.586790647813046]])
GOLDEN_NPV_2F = 1780380.7581833513    # test_multi_factor.py:226
GOLDEN_NPV_3F = 1766460.137569665     # test_multi_factor.py:298


def spot_sim_problem(num_sims):   # test_multi_factor.py:34-70
    S, F = 3, 3
    cur = pd.Timestamp("2020-07-27")
    dates = [pd.Timestamp("2020-08-01"), pd.Timestamp("2021-01-15"), pd.Timestamp("2021-07-30")]
    sim_time = np.array([(d - cur).days / 365.0 for d in dates])
    vol = np.array([[0.35, 0.15, 0.95], [0.29, 0.18, 0.92], [0.32, 0.21, 0.89]])
    corr = np.array([[1.0, 0.6, 0.3], [0.6, 1.0, 0.4], [0.3, 0.4, 1.0]])
    fwd = np.array([56.85, 59.08, 62.453])
    z, o = np.zeros(S + 1), np.ones(S + 1)
    pad = lambda a: np.concatenate([a, a[-1:]])   # noqa: E731  a terminal dummy period so that P = S + 1
    return _abi.Problem(
        n_periods=S + 1, n_factors=F, n_basis=1, n_sims=num_sims, first_period_is_current=0, must_be_empty_at_end=1,
        start_inventory=0.0, grid_spacing=1.0, sim_time=pad(sim_time) + np.concatenate([np.zeros(S), [1.0 / 365.0]]),
        mean_reversion=[0.0, 2.5, 16.2], factor_vol=np.vstack([vol, vol[-1:]]), factor_corr=corr,
        forward=pad(fwd), spot_power=[0], factor_power=np.zeros((1, F), dtype=np.int32),
        min_inventory=z, max_inventory=z, constraint_id=np.zeros(S + 1, dtype=np.int32), inject_cost=z, inject_cost_df=o,
        withdraw_cost=z, withdraw_cost_df=o, consumed_inject=z, consumed_withdraw=z, inventory_loss=z, inventory_cost=z,
        inventory_cost_df=o, settle_df=o, constraint_kind=[_abi.CONSTRAINT_CONSTANT], constraint_first=[0],
        constraint_count=[1], pillar_inventory=[0.0], pillar_min_rate=[0.0], pillar_max_rate=[0.0]).finalize()


def test_seeded_spot_prices_of_the_reference_simulator():
    """seed 12, no antithetic, 4 sims x 3 dates, three factors (mean reversions 0, 2.5, 16.2): the oracle's simulator model on the
    reference's draws.  Eleven of the twelve prices come out bit for bit, the twelfth one ulp off (Math.Exp vs det_exp)."""
    z, _ = mt19937_ref.path_normals(12, False, 4, 3, 3)
    zz = np.concatenate([z, np.zeros((1, 3, 4))])            # the dummy terminal period draws nothing
    spot, _ = oracle.simulate(spot_sim_problem(4), Rng.from_normals(np.ascontiguousarray(zz)))
    rel = np.abs(spot[:3] / GOLDEN_SPOT - 1.0)
    assert rel.max() < 4e-16
    assert int((spot[:3] == GOLDEN_SPOT).sum()) >= 11


def wrapper_test_inputs():   # test_multi_factor.py:167-196
    storage = CmdtyStorage('D', '2019-12-01', '2020-04-01', 1.23, 0.98, min_inventory=0.0, max_inventory=100000.0,
                           max_injection_rate=700.0, max_withdrawal_rate=700.0)
    idx = pd.PeriodIndex([pd.Period(d, freq='D') for d in ['2019-08-29', '2020-03-12', '2020-04-01']])
    fwd = pd.Series([23.87, 150.32, 150.32], idx).resample('D').ffill()
    rates = pd.Series(0.03, index=pd.period_range('2019-08-29', '2020-06-01', freq='D'))
    return storage, fwd, rates


def twentieth_of_next_month(period):
    return period.asfreq('M').asfreq('D', 'end') + 20


def two_factor_problem(num_sims=500):
    storage, fwd, rates = wrapper_test_inputs()
    curve = pd.period_range('2019-08-29', '2020-06-01', freq='D')
    factors = [(0.0, pd.Series(0.14, index=curve)), (16.2, pd.Series(1.15, index=curve))]
    return flatten_problem(storage, '2019-08-29', 0.0, fwd, rates, twentieth_of_next_month, factors, np.array([[1.0, 0.64], [0.64, 1.0]]),
                           num_sims, '1 + x0 + x0**2 + x1 + x1*x1', False, None, 100, 1e-12)


def three_factor_problem(num_sims=500):
    storage, fwd, rates = wrapper_test_inputs()
    factors, corrs = create_3_factor_season_params('D', 16.2, 1.15, 0.14, 0.18, utils.to_period('2019-08-29', 'D'), storage.end)
    basis = '1 + x_st + x_sw + x_lt + x_st**2 + x_sw**2 + x_lt**2'.replace('x_st', 'x0').replace('x_lt', 'x1').replace('x_sw', 'x2')
    return flatten_problem(storage, '2019-08-29', 0.0, fwd, rates, twentieth_of_next_month, factors, corrs, num_sims, basis, False, None, 100, 1e-12)


@pytest.mark.parametrize("make,golden", [(two_factor_problem, GOLDEN_NPV_2F), (three_factor_problem, GOLDEN_NPV_3F)])
def test_seeded_npv_of_the_reference_wrapper_tests_bit_for_bit(make, golden):
    """seed = fwd_sim_seed = 11, 500 sims, antithetic: both simulation sets are the same draws (two generators with one seed)."""
    pb = make()
    z, _ = mt19937_ref.path_normals(11, True, int(pb.n_sims), pb.n_sim_periods, pb.n_factors)
    res = oracle.calculate(pb, Rng.from_normals(z), Rng.from_normals(z)).results
    assert res.npv == golden   # the reference asserts 6 decimal places; the oracle gives every bit


def test_product_generator_equals_the_test_side_restatement():
    """csrc/mt19937.h (behind LSMC_RNG_MT19937 and lsmc_mt19937_normals) against tests/mt19937_ref.py: two independent
    implementations of the same stream, with and without antithetic, odd and even path counts."""
    for seed, anti, n, s, f in [(12, False, 4, 3, 3), (11, True, 500, 123, 2), (7, True, 7, 5, 3), (123456789, False, 9, 2, 1)]:
        got = _native.mt19937_normals(seed, anti, n, s, f)
        want, _ = mt19937_ref.path_normals(seed, anti, n, s, f)
        assert np.array_equal(got, want), (seed, anti, n, s, f)


# ---- the values the reference PRINTS for its ratchet storage (README.md:168-173, samples/python/multifactor_storage.ipynb cells 21-31) --------
def readme_inputs(injection_cost, withdrawal_cost):   # README.md:104-163 / notebook cells 4-19
    from storage_b200 import RatchetInterp
    storage = CmdtyStorage(freq='D', storage_start='2021-04-01', storage_end='2022-04-01', injection_cost=injection_cost, withdrawal_cost=withdrawal_cost,
                           ratchets=[('2021-04-01', [(0.0, -150.0, 250.0), (2000.0, -200.0, 175.0), (5000.0, -260.0, 155.0), (7000.0, -275.0, 132.0)]),
                                     ('2022-10-01', [(0.0, -130.0, 260.0), (2000.0, -190.0, 190.0), (5000.0, -230.0, 165.0), (7000.0, -245.0, 148.0)])],
                           ratchet_interp=RatchetInterp.LINEAR)
    prices = [16.61, 15.68, 15.42, 15.31, 15.27, 15.13, 15.96, 17.22, 17.32, 17.66, 17.59, 16.81, 15.36, 14.49, 14.28, 14.25, 14.32, 14.33, 15.30,
              16.58, 16.64, 16.79, 16.64, 15.90, 14.63]
    fwd = pd.Series(data=prices, index=pd.period_range(start='2021-04-25', periods=25, freq='M')).resample('D').ffill()
    pillars = pd.PeriodIndex(freq='D', data=['2021-04-25', '2021-06-01', '2021-08-01', '2021-12-01', '2022-04-01', '2022-12-01', '2023-12-01'])
    rates = pd.Series(data=[0.005, 0.006, 0.0072, 0.0087, 0.0101, 0.0115, 0.0126], index=pillars).resample('D').asfreq('D').interpolate(method='linear')
    return storage, fwd, rates


def test_values_printed_by_the_readme_and_the_notebook():
    """Ratchets, injection / withdrawal costs, discounted deltas, spot price among the ten basis functions, seeds 12 and 25, and a
    valuation set that CONTINUES the regression generator (README: no fwd_sim_seed).  The README's code block shows costs 0.01 / 0.025
    but its printed numbers are those of the notebook's 0.1 / 0.25 (the intrinsic value, which needs no random numbers, says so)."""
    from storage_b200 import intrinsic_value
    storage, fwd, rates = readme_inputs(0.1, 0.25)
    args = (storage, '2021-04-25', 1500.0, fwd, rates, twentieth_of_next_month)
    intrinsic = intrinsic_value(*args).npv
    assert '{0:,.0f}'.format(intrinsic) == '38,446'                                              # README.md:172, notebook cell 31
    assert '{0:,.0f}'.format(intrinsic_value(*((readme_inputs(0.01, 0.025)[0],) + args[1:])).npv) != '38,446'
    factors, corrs = create_3_factor_season_params('D', 91.0, 0.85, 0.30, 0.19, utils.to_period('2021-04-25', 'D'), storage.end)
    basis = '1 + x_st + x_sw + x_lt + s + x_st**2 + x_sw**2 + x_lt**2 + s**2 + s * x_st'.replace('x_st', 'x0').replace('x_lt', 'x1').replace('x_sw', 'x2')

    def lsmc(factors, corrs, basis, num_sims, seed, fwd_seed):
        pb = flatten_problem(*args, factors, corrs, num_sims, basis, True, None, 100, 1e-12)
        S, F = pb.n_sim_periods, pb.n_factors
        if fwd_seed is None:   # one generator, 2 x num_sims calls
            z = _native.mt19937_normals(seed, True, 2 * num_sims, S, F)
            zr, zv = np.ascontiguousarray(z[:, :, :num_sims]), np.ascontiguousarray(z[:, :, num_sims:])
        else:
            zr, zv = _native.mt19937_normals(seed, True, num_sims, S, F), _native.mt19937_normals(fwd_seed, True, num_sims, S, F)
        return oracle.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv)).results.npv

    notebook = lsmc(factors, corrs, basis, 500, 12, 25)
    assert '{0:,.0f}'.format(notebook) == '69,524' and '{0:,.0f}'.format(notebook - intrinsic) == '31,078'     # notebook cells 21, 31
    readme = lsmc(factors, corrs, basis, 2000, 12, None)
    assert '{0:,.0f}'.format(readme) == '69,496' and '{0:,.0f}'.format(readme - intrinsic) == '31,049'         # README.md:171-173
    vol = pd.Series(0.64, index=pd.period_range(start='2021-04-25', end=storage.end, freq='D'))
    two_factor = lsmc([(16.2, vol), (0.0, vol)], np.array([[1.0, 0.64], [0.64, 1.0]]),
                      '1 + x0 + x1 + s + x0**2 + x1**2 + s**2 + x0*x1 + x1*s + x0*s', 500, 12, 25)
    assert '{0:,.0f}'.format(two_factor) == '63,505'                                              # notebook cell 29
