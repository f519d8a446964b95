"""Host-side pieces of the Python entry points that need no GPU: the factor / correlation validator (messages are the contract:
src/Cmdty.Storage.Python/cmdty_storage/multi_factor.py:112-146) and the three-factor seasonal parameterisation (:254-289)."""
import math

import numpy as np
import pandas as pd
import pytest

from storage_b200.multi_factor import _validate_multi_factor_params
from storage_b200.multi_factor_params import create_3_factor_season_params


def _curve(n=3):
    return pd.Series([0.5] * n, index=pd.period_range("2021-01-01", periods=n, freq="D"))


def test_validator_fills_in_the_one_and_two_factor_shorthands():
    assert np.array_equal(_validate_multi_factor_params([(0.0, _curve())], None), [[1.0]])
    assert np.array_equal(_validate_multi_factor_params([(0.0, _curve()), (1.0, _curve())], 0.25), [[1.0, 0.25], [0.25, 1.0]])
    full = np.array([[1.0, 0.6, 0.3], [0.6, 1.0, 0.4], [0.3, 0.4, 1.0]])
    out = _validate_multi_factor_params([(0.0, _curve())] * 3, full)
    assert out.dtype == np.float64 and np.array_equal(out, full)


@pytest.mark.parametrize("n_factors, corrs, message", [
    (0, None, "factors cannot be empty."),
    (2, None, "Factor correlation matrix is not 2-dimensional."),
    (3, 0.5, "Factor correlation matrix is not 2-dimensional."),
    (2, np.ones((2, 3)), "Factor correlation matrix is not square."),
    (2, np.array([[1.0, 0.2], [0.2, 0.9]]), "Factor correlation on diagonal position (1, 1) value of 0.9 not valid as not equal to 1."),
    (2, np.array([[1.0, 1.5], [-1.2, 1.0]]), "Factor correlation in position (0, 1) value of 1.5 not valid as not in the interval [-1, 1]"),
    (2, np.array([[1.0, 0.5], [-1.2, 1.0]]), "Factor correlation in position (1, 0) value of -1.2 not valid as not in the interval [-1, 1]"),
    (2, np.array([[1.0, float("nan")], [0.1, 1.0]]), "Factor correlation in position (0, 1) value of nan not valid as not in the interval [-1, 1]"),
    (3, np.eye(2), "factor_corrs and factors must be of the same length. factor_corrs is of length 2 and factors is of length 3."),
])
def test_validator_messages(n_factors, corrs, message):
    with pytest.raises(ValueError) as e:
        _validate_multi_factor_params([(0.0, _curve())] * n_factors, corrs)
    assert str(e.value) == message


def test_first_offender_in_row_major_order_is_reported():
    bad = np.array([[1.0, 0.1, 0.1], [0.1, 1.2, 7.0], [9.0, 0.1, 1.0]])
    with pytest.raises(ValueError, match=r"diagonal position \(1, 1\)"):
        _validate_multi_factor_params([(0.0, _curve())] * 3, bad)


@pytest.mark.parametrize("freq, start, end", [("D", "2019-08-28", "2020-04-01"), ("H", "2021-01-30", "2021-02-02"),
                                              ("30min", "2021-01-31", "2021-02-01"), ("M", "2020-01", "2021-06")])
def test_three_factor_seasonal_parameterisation(freq, start, end):
    factors, corrs = create_3_factor_season_params(freq, 45.0, 1.2, 0.3, 0.19, start, end)
    assert np.array_equal(corrs, np.eye(3))
    (mr_spot, spot), (mr_long, long_term), (mr_season, season) = factors
    assert (mr_spot, mr_long, mr_season) == (45.0, 0.0, 0.0)
    assert spot.index.equals(long_term.index) and spot.index.equals(season.index)
    assert spot.index[0] == pd.Period(start, freq=spot.index.freq) and spot.index[-1] == pd.Period(end, freq=spot.index.freq)
    assert (spot == 1.2).all() and (long_term == 0.3).all()
    # the definition, computed independently: seconds from 1 Feb (start year) 00:00 as integer arithmetic on the timestamps
    feb1 = pd.Timestamp(year=spot.index[0].year, month=2, day=1)
    for p, v in season.items():
        seconds = (p.start_time.value - feb1.value) // 10 ** 9
        t = seconds / (365.25 * 86400.0)
        assert v == math.sin(2.0 * math.pi * t + math.pi / 2.0) * 0.095 or abs(v - math.cos(2.0 * math.pi * t) * 0.095) < 1e-15
    if freq == "D":   # peak on 1 February, trough half a year later
        assert season[pd.Period("2020-02-01", freq="D")] == pytest.approx(0.095 * math.cos(2 * math.pi * 365.0 / 365.25), abs=1e-15)
        assert season.max() <= 0.095 and season.min() >= -0.095


def test_build_config_grid_points_override_sets_the_grid_spacing():
    """configs.build_config(grid_points=n): the parity cases with inventory grids beyond the pipelined sweep's 104 rows."""
    from storage_b200.configs import build_config
    coarse = build_config("c3", num_sims=8, end="2021-06-01")
    fine = build_config("c3", num_sims=8, end="2021-06-01", grid_points=130)
    assert coarse.grid_spacing == pytest.approx(fine.grid_spacing * 129.0 / 99.0, rel=1e-12)
    assert fine.n_periods == coarse.n_periods and fine.n_basis == coarse.n_basis
