"""intrinsic_value — the reference's deterministic valuation on the forward curve
(src/Cmdty.Storage.Python/cmdty_storage/intrinsic.py:42-85 -> IntrinsicStorageValuation.cs:120-322) with the same signature
and result tuple, computed by the host routine lsmc_intrinsic of liblsmc_b200.so (include/lsmc_b200.h).  The multi-factor
entry points call it beside the LSMC valuation, as the reference does (multi_factor.py:406), to fill intrinsic_npv /
intrinsic_profile / extrinsic_npv.  It is O(periods x grid points) host arithmetic and needs no GPU."""
import typing as tp
from datetime import date

import numpy as np
import pandas as pd

from . import _native, utils
from .cmdty_storage import CmdtyStorage
from .flatten import flatten_problem, TrivialResult

PROFILE_COLUMNS = ['inventory', 'inject_withdraw_volume', 'cmdty_consumed', 'inventory_loss', 'net_volume', 'period_pv']


class IntrinsicValuationResults(tp.NamedTuple):
    npv: float
    profile: pd.DataFrame


def intrinsic_of_problem(problem) -> IntrinsicValuationResults:
    """Intrinsic value of an already flattened problem (or of a trivial exit, IntrinsicStorageValuation.cs:128-151)."""
    if isinstance(problem, TrivialResult):
        return IntrinsicValuationResults(problem.npv, pd.DataFrame(columns=PROFILE_COLUMNS, dtype=np.float64))
    npv, profile = _native.intrinsic(problem)
    return IntrinsicValuationResults(npv, pd.DataFrame(profile, index=problem.periods, columns=PROFILE_COLUMNS))


def intrinsic_value(cmdty_storage: CmdtyStorage,
                    val_date: utils.TimePeriodSpecType,
                    inventory: tp.Union[float, int],
                    forward_curve: pd.Series,
                    interest_rates: pd.Series,
                    settlement_rule: tp.Callable[[pd.Period], date],
                    num_inventory_grid_points: int = 100,
                    numerical_tolerance: float = 1E-12) -> IntrinsicValuationResults:
    if utils.pandas_freq(cmdty_storage.freq) != utils.pandas_freq_of_index(forward_curve.index):
        raise ValueError("cmdty_storage and forward_curve have different frequencies.")
    # the price model plays no part: one placeholder factor and a constant basis complete the problem description
    vol = pd.Series(1.0, index=forward_curve.index)
    problem = flatten_problem(cmdty_storage, val_date, inventory, forward_curve, interest_rates, settlement_rule, [(0.0, vol)],
                              np.array([[1.0]]), 1, '1', False, None, num_inventory_grid_points, numerical_tolerance)
    return intrinsic_of_problem(problem)
