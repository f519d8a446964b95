"""Host-side helpers mirroring cmdty_storage/utils.py and time_func.py of the reference
(src/Cmdty.Storage.Python/cmdty_storage/utils.py:47-135, time_func.py:30-42) without pythonnet."""
import re
import typing as tp
from datetime import date, datetime

import dateutil.parser
import numpy as np
import pandas as pd

TimePeriodSpecType = tp.Union[datetime, date, str, pd.Period]
ForwardPointType = tp.Union[datetime, date, str, pd.Period]
CurveType = tp.Union[pd.Series, tp.Dict[ForwardPointType, float]]

# Reference keys (utils.py:120-127) -> pandas period alias understood by the installed pandas.
FREQ_TO_PERIOD_TYPE = {"15min": "15min", "30min": "30min", "H": "h", "D": "D", "M": "M", "Q": "Q"}


def pandas_freq(freq: str) -> str:
    if freq not in FREQ_TO_PERIOD_TYPE:
        raise ValueError("freq parameter value of '{}' not supported. The allowable values can be found in the keys of "
                         "the dict curves.FREQ_TO_PERIOD_TYPE.".format(freq))
    return pd.Period("2000-01-01", freq=FREQ_TO_PERIOD_TYPE[freq]).freqstr


def pandas_freq_of_index(index: pd.PeriodIndex) -> str:
    """Canonical pandas alias of a PeriodIndex's frequency (the reference compares `freqstr` with the storage freq)."""
    return pd.Period("2000-01-01", freq=index.freq).freqstr


def to_period(date_like: TimePeriodSpecType, freq: str) -> pd.Period:
    """utils.from_datetime_like (utils.py:47-50): any date-like -> the period of `freq` containing it."""
    pf = pandas_freq(freq)
    if isinstance(date_like, pd.Period):
        return date_like.asfreq(pf, "s") if date_like.freqstr != pd.Period("2000-01-01", freq=pf).freqstr else date_like
    if isinstance(date_like, str):
        date_like = dateutil.parser.parse(date_like)
    return pd.Period(date_like, freq=pf)


def first_day(period: pd.Period) -> pd.Period:
    """period.First<Day>()"""
    return period.asfreq("D", "s")


def to_day(date_like) -> pd.Period:
    if isinstance(date_like, pd.Period):
        return date_like.asfreq("D", "s")
    if isinstance(date_like, str):
        date_like = dateutil.parser.parse(date_like)
    return pd.Period(date_like, freq="D")


def act_365(start, end) -> float:
    """time_func.act_365 (time_func.py:30-33): whole days / 365."""
    return (to_day(end).ordinal - to_day(start).ordinal) / 365


def act_365_datetime(start: pd.Timestamp, end: pd.Timestamp) -> float:
    """Cmdty.Core TimeFunctions.Act365(DateTime, DateTime) as used at LVP.cs:176: fractional days / 365."""
    return (end - start).total_seconds() / 86400.0 / 365.0


def is_scalar(arg) -> bool:
    return isinstance(arg, (int, float, np.integer, np.floating)) and not isinstance(arg, bool)


def raise_if_none(arg, error_message):
    if arg is None:
        raise ValueError(error_message)


def raise_if_not_none(arg, error_message):
    if arg is not None:
        raise ValueError(error_message)


def curve_lookup(curve: CurveType, period: pd.Period, freq: str, what: str) -> float:
    """TimeSeries<T,double>[period] / curve_to_net_dict lookup: exact period match or raise."""
    if isinstance(curve, pd.Series):
        try:
            return float(curve[period])
        except KeyError:
            raise ValueError("No {} provided for {}.".format(what, period))
    for key, value in curve.items():
        if to_period(key, freq) == period:
            return float(value)
    raise ValueError("No {} provided for {}.".format(what, period))


def series_values_for(series: pd.Series, periods: pd.PeriodIndex, what: str) -> np.ndarray:
    """Vectorised `series[period]` for all `periods`; raises like the reference when a period is missing."""
    if not isinstance(series.index, pd.PeriodIndex):
        series = pd.Series(series.values, index=pd.PeriodIndex(series.index, freq=periods.freq))
    idx = series.index.get_indexer(periods)
    if (idx < 0).any():
        missing = periods[int(np.argmax(idx < 0))]
        raise ValueError("No {} provided for {}.".format(what, missing))
    return series.values[idx].astype(np.float64)


_MONOMIAL_TERM = re.compile(r"^(s|x(\d+)|1)(?:\*\*(\d+))?$")


def parse_basis_functions(expression: str, num_factors: int) -> tp.Tuple[np.ndarray, np.ndarray]:
    """BasisFunctionsBuilder.Parse (BasisFunctionsBuilder.cs:90-103,116-130) without Roslyn.

    Grammar: monomials joined by '+', each a product ('*') of `1`, `s`, `xN`, optionally raised with `**q`.
    Returns (spot_power[K], factor_power[K][F]) — the exponents PowerMonomialBuilder hides inside its delegate
    (PowerMonomialBuilder.cs:31-70, Sim.cs:30-37).
    """
    if expression is None:
        raise ValueError("basis function expression is None")
    monomials = [m.strip() for m in expression.split("+")]
    if len(monomials) == 0 or any(m == "" for m in monomials):
        raise ValueError("Basis function expression contains no monomials.")
    if len(set(monomials)) < len(monomials):
        raise ValueError("Basis function expression contains repeated monomials.")
    spot_power = np.zeros(len(monomials), dtype=np.int32)
    factor_power = np.zeros((len(monomials), num_factors), dtype=np.int32)
    for k, mono in enumerate(monomials):
        text = mono.replace(" ", "").replace("S", "s")
        # split on single '*' but not on '**'
        terms = re.split(r"(?<!\*)\*(?!\*)", text)
        for term in terms:
            m = _MONOMIAL_TERM.match(term)
            if not m:
                raise ValueError("Cannot parse basis function monomial '{}'.".format(mono))
            power = int(m.group(3)) if m.group(3) is not None else 1
            if m.group(1) == "1":
                continue
            if m.group(1) == "s":
                spot_power[k] += power
            else:
                f = int(m.group(2))
                if f >= num_factors:
                    raise ValueError("Basis function refers to factor x{} but the model has {} factors.".format(f, num_factors))
                factor_power[k, f] += power
    return spot_power, factor_power
