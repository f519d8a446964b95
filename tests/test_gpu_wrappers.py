"""The reference's Python entry points for the LSMC path, reproduced by storage_b200 with the same names and arguments
(multi_factor_value, three_factor_seasonal_value, MultiFactorSpotSim; src/Cmdty.Storage.Python/cmdty_storage/
multi_factor.py:49-92,324-455), run on the GPU and held to the reference's own wrapper tests
(src/Cmdty.Storage.Python/tests/test_multi_factor.py:167-311): result shapes, 123 deltas, 245 progress calls, the golden
intrinsic NPV, and the SEEDED LSMC NPVs 1780380.7581833513 / 1766460.137569665 — seeded calls draw the reference's own stream
(LSMC_RNG_MT19937: MathNet MT19937 + Normal.Samples in Cmdty.Core's draw order, csrc/mt19937.h), which the CPU oracle turns into
those numbers bit for bit (tests/test_reference_goldens.py) and the device path to 1e-9 or better (its regression solves
preconditioned normal equations instead of a Householder QR).  The twelve seeded spot prices of the simulator test too."""
import numpy as np
import pandas as pd
import pytest

import storage_b200
from storage_b200 import CmdtyStorage, multi_factor
from storage_b200._abi import Rng
from storage_b200.flatten import flatten_problem
from oracle import oracle
from test_intrinsic import wrapper_test_inputs, twentieth_of_next_month

pytestmark = pytest.mark.gpu
VAL_DATE, NUM_SIMS, SEED = '2019-08-29', 500, 11


def check_result_structure(res, progresses):
    assert len(res.deltas) == 123 and len(res.expected_profile) == 123            # test_multi_factor.py:227-228
    assert progresses[-1] == 1.0 and len(progresses) == 245                       # :229-230
    assert all(b > a for a, b in zip(progresses, progresses[1:]))
    assert res.intrinsic_npv == 1703773.0757192627 and len(res.intrinsic_profile) == 123   # :231-232
    assert res.extrinsic_npv == res.npv - res.intrinsic_npv and res.extrinsic_npv > 0.0
    for frame in (res.sim_spot_regress, res.sim_spot_valuation, res.sim_inventory, res.sim_inject_withdraw, res.sim_cmdty_consumed,
                  res.sim_inventory_loss, res.sim_net_volume, res.sim_pv):
        assert frame.shape == (123, NUM_SIMS)                                     # :233-239
    assert list(res.expected_profile.columns) == ['inventory', 'inject_withdraw_volume', 'cmdty_consumed', 'inventory_loss', 'net_volume', 'period_pv']
    assert list(res.trigger_prices.columns) == ['inject_volume', 'inject_trigger_price', 'withdraw_volume', 'withdraw_trigger_price']
    assert len(res.trigger_prices) == 122 and len(res.trigger_profiles) == 122
    assert res.deltas.index[0] == pd.Period('2019-12-01', freq='D') and res.deltas.iloc[-1] == 0.0
    assert res.sim_pv.sum(axis=0).mean() == pytest.approx(res.npv, rel=1e-12)     # NPV = mean over sims of the summed period PVs
    assert res.expected_profile['inventory'].iloc[0] == 0.0


def test_multi_factor_value_regression():   # test_multi_factor.py:167-239
    storage, fwd, rates = wrapper_test_inputs()
    idx = pd.period_range(VAL_DATE, '2020-06-01', freq='D')
    factors = [(0.0, pd.Series(0.14, index=idx)), (16.2, pd.Series(1.15, index=idx))]
    progresses = []
    res = storage_b200.multi_factor_value(storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, factors, 0.64, NUM_SIMS,
                                          '1 + x0 + x0**2 + x1 + x1*x1', False, seed=SEED, fwd_sim_seed=SEED,
                                          on_progress_update=progresses.append)
    check_result_structure(res, progresses)
    print("multi_factor_value: npv %.10f, relative to the reference's golden value %.3e" % (res.npv, res.npv / 1780380.7581833513 - 1.0))
    assert round(res.npv - 1780380.7581833513, 6) == 0           # the reference's own assertion (:226, assertAlmostEqual places=6)
    # the same flattened problem and draws through the CPU oracle
    pb = flatten_problem(storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, factors, np.array([[1.0, 0.64], [0.64, 1.0]]),
                         NUM_SIMS, '1 + x0 + x0**2 + x1 + x1*x1', False)
    z = storage_b200._native.mt19937_normals(SEED, True, NUM_SIMS, pb.n_sim_periods, pb.n_factors)
    ref = oracle.calculate(pb, Rng.from_normals(z), Rng.from_normals(z)).results
    assert ref.npv == 1780380.7581833513
    assert abs(res.npv / ref.npv - 1.0) < 1e-9
    assert np.max(np.abs(res.deltas.values - ref.deltas)) < 1e-9 * np.max(np.abs(ref.deltas))
    assert np.max(np.abs(res.expected_profile.values - ref.profile)) < 1e-9 * np.max(np.abs(ref.profile))


def test_three_factor_seasonal_regression():   # test_multi_factor.py:241-311
    storage, fwd, rates = wrapper_test_inputs()
    progresses = []
    res = storage_b200.three_factor_seasonal_value(storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, 16.2, 1.15, 0.14, 0.18,
                                                   NUM_SIMS, '1 + x_st + x_sw + x_lt + x_st**2 + x_sw**2 + x_lt**2', False,
                                                   seed=SEED, fwd_sim_seed=SEED, on_progress_update=progresses.append)
    check_result_structure(res, progresses)
    print("three_factor_seasonal_value: npv %.10f, relative to the reference's golden value %.3e" % (res.npv, res.npv / 1766460.137569665 - 1.0))
    assert round(res.npv - 1766460.137569665, 6) == 0            # the reference's own assertion (:298, places=6)


def test_seeded_runs_continue_the_regression_generator_when_no_forward_seed_is_given():
    """LVP.cs:186-187: valuationSimSeed == null -> the regression generator continues its sequence for the valuation set
    (LSMC_RNG_MT19937_CONTINUE), so the two sets differ and the NPV differs from the fwd_sim_seed = seed run; Philox stays
    available for seeded runs with very many paths (multi_factor.SEEDED_RNG)."""
    storage, fwd, rates = wrapper_test_inputs()
    idx = pd.period_range(VAL_DATE, '2020-06-01', freq='D')
    factors = [(0.0, pd.Series(0.14, index=idx)), (16.2, pd.Series(1.15, index=idx))]
    args = (storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, factors, 0.64, NUM_SIMS, '1 + x0 + x0**2 + x1 + x1*x1', False)
    cont = storage_b200.multi_factor_value(*args, seed=SEED)
    pb = flatten_problem(storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, factors, np.array([[1.0, 0.64], [0.64, 1.0]]),
                         NUM_SIMS, '1 + x0 + x0**2 + x1 + x1*x1', False)
    z = storage_b200._native.mt19937_normals(SEED, True, 2 * NUM_SIMS, pb.n_sim_periods, pb.n_factors)   # one generator, 2 x 500 calls
    ref = oracle.calculate(pb, Rng.from_normals(np.ascontiguousarray(z[:, :, :NUM_SIMS])), Rng.from_normals(np.ascontiguousarray(z[:, :, NUM_SIMS:]))).results
    assert abs(cont.npv / ref.npv - 1.0) < 1e-9
    assert not np.array_equal(cont.sim_spot_regress.values, cont.sim_spot_valuation.values)
    assert abs(cont.npv / 1780380.7581833513 - 1.0) < 0.03 and cont.npv != 1780380.7581833513
    multi_factor.SEEDED_RNG = "philox"
    try:
        dev = storage_b200.multi_factor_value(*args, seed=SEED, fwd_sim_seed=SEED)
    finally:
        multi_factor.SEEDED_RNG = "mt19937"
    assert abs(dev.npv / 1780380.7581833513 - 1.0) < 0.03 and abs(dev.npv / 1780380.7581833513 - 1.0) > 1e-9


def test_seeded_spot_prices_of_the_reference_simulator_on_the_device():   # test_multi_factor.py:34-91
    from datetime import date
    factors = [(0.0, {date(2020, 8, 1): 0.35, '2021-01-15': 0.29, date(2021, 7, 30): 0.32}),
               (2.5, pd.Series(data=[0.15, 0.18, 0.21], index=pd.PeriodIndex(data=['2020-08-01', '2021-01-15', '2021-07-30'], freq='D'))),
               (16.2, {date(2020, 8, 1): 0.95, '2021-01-15': 0.92, date(2021, 7, 30): 0.89})]
    corrs = np.array([[1.0, 0.6, 0.3], [0.6, 1.0, 0.4], [0.3, 0.4, 1.0]])
    fwd_curve = {'2020-08-01': 56.85, pd.Period('2021-01-15', freq='D'): 59.08, date(2021, 7, 30): 62.453}
    sim = multi_factor.MultiFactorSpotSim('D', factors, corrs, date(2020, 7, 27), fwd_curve,
                                          [pd.Period('2020-08-01'), '2021-01-15', date(2021, 7, 30)], 12)
    spots = sim.simulate(4)
    golden = np.array([[52.59976397688973, 46.1206448628463, 58.15838580682589, 65.500441945042979],
                       [57.559631642935514, 72.0381089486175, 82.49607173562342, 42.812676607997183],
                       [89.40526992772634, 85.18869803117379, 138.68587285875978, 76.586790647813046]])
    assert spots.shape == (3, 4)
    assert np.max(np.abs(spots.values / golden - 1.0)) < 4e-16 and int((spots.values == golden).sum()) >= 11


def test_wrapper_validates_like_the_reference():
    storage, fwd, rates = wrapper_test_inputs()
    idx = pd.period_range(VAL_DATE, '2020-06-01', freq='D')
    vol = pd.Series(0.5, index=idx)
    with pytest.raises(ValueError, match="factors cannot be empty"):                               # multi_factor.py:113-115
        storage_b200.multi_factor_value(storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, [], None, 10, '1 + x0', False)
    with pytest.raises(ValueError, match="not valid as not in the interval"):                      # :137-141
        storage_b200.multi_factor_value(storage, VAL_DATE, 0.0, fwd, rates, twentieth_of_next_month, [(1.0, vol), (2.0, vol)], 1.5, 10,
                                        '1 + x0', False)
    monthly = pd.Series(50.0, index=pd.period_range('2019-08', '2020-06', freq='M'))
    with pytest.raises(ValueError, match="different frequencies"):                                 # :372-373
        storage_b200.multi_factor_value(storage, VAL_DATE, 0.0, monthly, rates, twentieth_of_next_month, [(1.0, vol)], None, 10,
                                        '1 + x0', False)
    with pytest.raises(ValueError, match="Inventory cannot be negative"):                          # LSV.cs:61
        storage_b200.multi_factor_value(storage, VAL_DATE, -1.0, fwd, rates, twentieth_of_next_month, [(1.0, vol)], None, 10, '1 + x0', False)
    expired = storage_b200.multi_factor_value(storage, '2020-04-05', 0.0, pd.Series(50.0, index=pd.period_range('2020-04-01', '2020-05-01', freq='D')),
                                              rates, twentieth_of_next_month, [(1.0, vol)], None, 10, '1 + x0', False)
    assert expired.npv == 0.0 and len(expired.deltas) == 0                                         # LSV.cs:64-68


def test_multi_factor_spot_sim_shapes_mean_and_variance():   # multi_factor.py:49-92; test_multi_factor.py:45-91
    idx = pd.period_range('2020-08-01', '2021-01-01', freq='D')
    fwd = pd.Series(50.0 + 0.01 * np.arange(len(idx)), index=idx)
    vol = pd.Series(0.6, index=idx)
    sim_periods = pd.period_range('2020-09-01', '2020-09-30', freq='D')
    sim = multi_factor.MultiFactorSpotSim('D', [(12.0, vol)], None, '2020-08-01', fwd, sim_periods, seed=12, antithetic=True)
    n = 20_000
    spots = sim.simulate(n)
    assert spots.shape == (30, n) and list(spots.index) == list(sim_periods)
    mean = spots.mean(axis=1).values
    se = spots.std(axis=1).values / np.sqrt(n / 2)
    assert np.all(np.abs(mean - fwd[sim_periods].values) < 4.5 * se)                 # martingale: E[S(T)] = F(0,T)
    t = np.array([(p.start_time - pd.Timestamp('2020-08-01')).days / 365.0 for p in sim_periods])
    var = 0.6 ** 2 * (1.0 - np.exp(-2 * 12.0 * t)) / (2 * 12.0)                      # TestHelper.cs:77-86
    assert np.allclose(np.log(spots.values).var(axis=1), var, rtol=0.05)
    again = multi_factor.MultiFactorSpotSim('D', [(12.0, vol)], None, '2020-08-01', fwd, sim_periods, seed=12, antithetic=True).simulate(n)
    assert np.array_equal(again.values, spots.values)                                # seeded: reproducible


def test_readme_and_notebook_values_through_the_wrappers_on_the_device():
    """README.md:137-173 and samples/python/multifactor_storage.ipynb cells 21-31, called exactly as written there (the printed numbers
    are those of the notebook's injection / withdrawal costs 0.1 / 0.25: tests/test_reference_goldens.py)."""
    from test_reference_goldens import readme_inputs
    storage, fwd, rates = readme_inputs(0.1, 0.25)

    def settlement_rule(delivery_date):
        return delivery_date.asfreq('M').asfreq('D', 'end') + 20
    basis = '1 + x_st + x_sw + x_lt + s + x_st**2 + x_sw**2 + x_lt**2 + s**2 + s * x_st'
    readme = storage_b200.three_factor_seasonal_value(cmdty_storage=storage, val_date='2021-04-25', inventory=1500.0, fwd_curve=fwd,
                                                      interest_rates=rates, settlement_rule=settlement_rule, num_sims=2000, seed=12,
                                                      spot_mean_reversion=91.0, spot_vol=0.85, long_term_vol=0.30, seasonal_vol=0.19,
                                                      basis_funcs=basis, discount_deltas=True)
    assert '{0:,.0f}'.format(readme.npv) == '69,496'                # README.md:171
    assert '{0:,.0f}'.format(readme.intrinsic_npv) == '38,446'      # :172
    assert '{0:,.0f}'.format(readme.extrinsic_npv) == '31,049'      # :173
    notebook = storage_b200.three_factor_seasonal_value(cmdty_storage=storage, val_date='2021-04-25', inventory=1500.0, fwd_curve=fwd,
                                                        interest_rates=rates, settlement_rule=settlement_rule, discount_deltas=True,
                                                        num_sims=500, seed=12, fwd_sim_seed=25, spot_mean_reversion=91.0, spot_vol=0.85,
                                                        long_term_vol=0.30, seasonal_vol=0.19, basis_funcs=basis)
    assert '{0:,.0f}'.format(notebook.npv) == '69,524' and '{0:,.0f}'.format(notebook.extrinsic_npv) == '31,078'   # notebook cells 21, 31
