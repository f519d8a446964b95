"""lsmc_intrinsic (the host routine behind storage_b200.intrinsic_value and the intrinsic_npv / intrinsic_profile fields of the
multi-factor results) against the reference's own vectors: the golden NPV its Python wrapper tests assert to the last
digit (src/Cmdty.Storage.Python/tests/test_multi_factor.py:231,303) and the known-answer tests of
tests/Cmdty.Storage.Test/IntrinsicStorageValuationTest.cs.  Host arithmetic only: runs without a GPU.

The golden value needs NO random numbers, so it pins — bit for bit — everything the LSMC path shares with the intrinsic
model: flattening of the storage and curves (storage_b200/flatten.py), Act/365 discount factors, the inventory space
(StorageHelper.CalculateInventorySpace), the fixed-spacing grids, the bang-bang decision sets and the cost arithmetic of
csrc/host_plan.cpp."""
import numpy as np
import pandas as pd
import pytest

from storage_b200 import CmdtyStorage, intrinsic_value, _native
from storage_b200._abi import Problem, CONSTRAINT_CONSTANT


def create_piecewise_flat_series(data, dt_index, freq):   # src/Cmdty.Storage.Python/tests/utils.py
    period_index = pd.PeriodIndex([pd.Period(dt, freq=freq) for dt in dt_index])
    return pd.Series(data, period_index).resample(freq).ffill()


def twentieth_of_next_month(period):
    return period.asfreq('M').asfreq('D', 'end') + 20


def wrapper_test_inputs():   # test_multi_factor.py:167-190
    storage = CmdtyStorage('D', '2019-12-01', '2020-04-01', 1.23, 0.98, min_inventory=0.0, max_inventory=100000.0,
                           max_injection_rate=700.0, max_withdrawal_rate=700.0)
    fwd = create_piecewise_flat_series([23.87, 150.32, 150.32], ['2019-08-29', '2020-03-12', '2020-04-01'], freq='D')
    rates = pd.Series(0.03, index=pd.period_range('2019-08-29', '2020-06-01', freq='D'))
    return storage, fwd, rates


def test_reference_golden_intrinsic_npv_bit_for_bit():
    storage, fwd, rates = wrapper_test_inputs()
    res = intrinsic_value(storage, '2019-08-29', 0.0, fwd, rates, twentieth_of_next_month)
    assert res.npv == 1703773.0757192627                    # test_multi_factor.py:231 assertEqual, not assertAlmostEqual
    assert len(res.profile) == 123                          # test_multi_factor.py:232
    assert list(res.profile.columns) == ['inventory', 'inject_withdraw_volume', 'cmdty_consumed', 'inventory_loss', 'net_volume', 'period_pv']
    assert res.profile['period_pv'].sum() == pytest.approx(res.npv, rel=1e-15)
    assert res.profile['inventory'].iloc[-1] == 0.0         # must be empty at the end


def september_problem(forward, current_day, start_inventory=0.0):
    """GenerateValuationResults (IntrinsicStorageValuationTest.cs:38-75): constant rates -45.5 / 56.6, inventory 0..1000, unit
    costs 0.8 / 1.2 paid on the day, no discounting, grid spacing 10, must be empty at the end of 2019-09-30."""
    periods = pd.period_range(f'2019-09-{current_day:02d}', '2019-09-30', freq='D')
    P = len(periods)
    z, o = np.zeros(P), np.ones(P)
    return Problem(
        n_periods=P, n_factors=1, n_basis=1, n_sims=1, first_period_is_current=1, must_be_empty_at_end=1,
        numerical_tolerance=1e-10, start_inventory=start_inventory, grid_spacing=10.0,
        sim_time=np.arange(P) / 365.0, mean_reversion=[0.0], factor_vol=np.ones((P, 1)), factor_corr=[[1.0]],
        forward=np.asarray(forward, dtype=np.float64), spot_power=[0], factor_power=np.zeros((1, 1), dtype=np.int32),
        min_inventory=z, max_inventory=np.concatenate([np.full(P - 1, 1000.0), [0.0]]), constraint_id=np.zeros(P, dtype=np.int32),
        inject_cost=np.full(P, 0.8), inject_cost_df=o, withdraw_cost=np.full(P, 1.2), withdraw_cost_df=o, consumed_inject=z,
        consumed_withdraw=z, inventory_loss=z, inventory_cost=z, inventory_cost_df=o, settle_df=o,
        constraint_kind=[CONSTRAINT_CONSTANT], constraint_first=[0], constraint_count=[1], pillar_inventory=[0.0],
        pillar_min_rate=[-45.5], pillar_max_rate=[56.6], periods=periods).finalize()


def test_zero_inventory_backwardated_curve_is_worth_nothing_and_does_nothing():   # IntrinsicStorageValuationTest.cs:103-128
    forward = 59.95 - 0.58 * np.arange(14, 30)      # GenerateBackwardatedCurve from 2019-09-01, read from the 15th on
    npv, profile = _native.intrinsic(september_problem(forward, 15))
    assert npv == 0.0
    assert np.all(profile == 0.0) and profile.shape == (16, 6)


@pytest.mark.parametrize("spread,expected", [(1.99, 0.0), (2.01, 7 * 45.5 * 0.01)])   # :131-183 (cycle cost 0.8 + 1.2)
def test_forward_spread_against_cycle_cost(spread, expected):
    forward = np.concatenate([np.full(8, 56.6), np.full(8, 56.6 + spread)])
    npv, profile = _native.intrinsic(september_problem(forward, 15))
    if expected == 0.0:
        assert npv == 0.0 and np.all(profile == 0.0)
    else:
        assert npv == pytest.approx(expected, abs=1e-10)    # Assert.Equal(expected, npv, 10)
        assert profile[:, 1].max() == pytest.approx(56.6, rel=1e-12) and profile[:, 1].min() == -45.5
        assert profile[:, 1].sum() == pytest.approx(0.0, abs=1e-9)          # what went in came out


def test_intrinsic_value_barely_depends_on_the_inventory_grid():
    """The value function is piecewise linear in the inventory: 50 and 400 grid points agree to a fraction of a percent."""
    storage, fwd, rates = wrapper_test_inputs()
    res = intrinsic_value(storage, '2019-08-29', 0.0, fwd, rates, twentieth_of_next_month, num_inventory_grid_points=50)
    fine = intrinsic_value(storage, '2019-08-29', 0.0, fwd, rates, twentieth_of_next_month, num_inventory_grid_points=400)
    assert abs(res.npv / fine.npv - 1.0) < 2e-3


def test_expired_storage_has_an_empty_profile():   # IntrinsicStorageValuationTest.cs:186-204, test_intrinsic.py:97-101
    storage, fwd, rates = wrapper_test_inputs()
    late_fwd = pd.Series(50.0, index=pd.period_range('2020-04-01', '2020-05-01', freq='D'))
    res = intrinsic_value(storage, '2020-04-02', 0.0, late_fwd, rates, twentieth_of_next_month)
    assert res.npv == 0.0 and len(res.profile) == 0
