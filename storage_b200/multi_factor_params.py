"""Multi-factor model parameter helpers (pure host code, no native dependency).

Mirrors src/Cmdty.Storage.Python/cmdty_storage/multi_factor.py:254-289."""
import typing as tp

import numpy as np
import pandas as pd

from . import utils

days_per_year = 365.25
seconds_per_year = 60 * 60 * 24 * days_per_year


def create_3_factor_season_params(freq: str, spot_mean_reversion: float, spot_vol: float, long_term_vol: float,
                                  seasonal_vol: float, start, end) -> tp.Tuple[tp.List[tp.Tuple[float, pd.Series]], np.ndarray]:
    """multi_factor.py:258-289: factors (spot: mean reverting, flat vol), (long-term: flat), (seasonal: sine vol curve
    peaking on 1 Feb), identity correlation."""
    factor_corrs = np.eye(3)
    pfreq = utils.pandas_freq(freq)
    start_period = start if isinstance(start, pd.Period) else pd.Period(start, freq=pfreq)
    end_period = end if isinstance(end, pd.Period) else pd.Period(end, freq=pfreq)
    index = pd.period_range(start=start_period, end=end_period, freq=start_period.freq)
    long_term_vol_curve = pd.Series(index=index, data=[long_term_vol] * len(index))
    spot_vol_curve = pd.Series(index=index.copy(), data=[spot_vol] * len(index))
    peak_period = pd.Period(year=start_period.year, month=2, day=1, freq=start_period.freq)
    phase = np.pi / 2.0
    amplitude = seasonal_vol / 2.0
    t_from_peak = np.array([(p.start_time - peak_period.start_time).total_seconds() / seconds_per_year for p in index])
    seasonal_vol_curve = pd.Series(index=index.copy(), data=np.sin(2.0 * np.pi * t_from_peak + phase) * amplitude)
    factors = [(spot_mean_reversion, spot_vol_curve), (0.0, long_term_vol_curve), (0.0, seasonal_vol_curve)]
    return factors, factor_corrs
