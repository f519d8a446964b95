"""The three-factor seasonal parameterisation (pure host code, no native dependency).

What it has to equal is the reference's definition, src/Cmdty.Storage.Python/cmdty_storage/multi_factor.py:254-289: a mean-reverting
spot factor and a long-term factor with flat volatility curves, a seasonal factor whose volatility curve is a sine of period one
year (365.25 days) with amplitude seasonal_vol / 2 that peaks on 1 February of the start year, and no correlation between them.
tests/test_wrapper_host_logic.py holds it against values computed from that definition."""
import typing as tp

import numpy as np
import pandas as pd

from . import utils

days_per_year = 365.25
seconds_per_year = 60 * 60 * 24 * days_per_year

FactorList = tp.List[tp.Tuple[float, pd.Series]]


def _season_curve(periods: pd.PeriodIndex, seasonal_vol: float) -> pd.Series:
    """(seasonal_vol / 2) sin(2 pi t + pi / 2), t = years from the start of 1 February of the first period's year to the start
    of the period."""
    first = periods[0]
    peak_start = pd.Period(year=first.year, month=2, day=1, freq=first.freq).start_time
    years = np.array([(p.start_time - peak_start).total_seconds() / seconds_per_year for p in periods])
    return pd.Series(np.sin(2.0 * np.pi * years + np.pi / 2.0) * (seasonal_vol / 2.0), index=periods.copy())


def create_3_factor_season_params(freq: str, spot_mean_reversion: float, spot_vol: float, long_term_vol: float,
                                  seasonal_vol: float, start, end) -> tp.Tuple[FactorList, np.ndarray]:
    pfreq = utils.pandas_freq(freq)
    first, last = (v if isinstance(v, pd.Period) else pd.Period(v, freq=pfreq) for v in (start, end))
    periods = pd.period_range(start=first, end=last, freq=first.freq)

    def flat(level):
        return pd.Series([level] * len(periods), index=periods.copy())

    factors = [(spot_mean_reversion, flat(spot_vol)), (0.0, flat(long_term_vol)), (0.0, _season_curve(periods, seasonal_vol))]
    return factors, np.eye(3)
