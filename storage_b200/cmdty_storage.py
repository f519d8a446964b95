"""CmdtyStorage — host-side mirror of cmdty_storage.CmdtyStorage
(src/Cmdty.Storage.Python/cmdty_storage/cmdty_storage.py:58-199) holding the storage as a parametric
description instead of a .NET object, so it can be flattened into the C-ABI tables (include/lsmc_b200.h).

Semantics follow the C# builder: CmdtyStorage.cs:173-569 and CmdtyStorageBuilderExtensions.cs:142-257.
"""
import bisect
import logging
import typing as tp
from enum import Enum

import numpy as np
import pandas as pd

from . import utils
from ._abi import CONSTRAINT_CONSTANT, CONSTRAINT_PIECEWISE_LINEAR, CONSTRAINT_STEP, CONSTRAINT_POLYNOMIAL

logger = logging.getLogger("cmdty.storage")


class InjectWithdrawRange(tp.NamedTuple):
    min_inject_withdraw_rate: float
    max_inject_withdraw_rate: float


class RatchetInterp(Enum):
    LINEAR = 1
    STEP = 2
    # Not in the reference's Python enum (cmdty_storage.py:44-46); the C# builder has it as InterpolationType.Polynomial
    # (CmdtyStorageBuilderExtensions.cs:78-82 -> PolynomialInjectWithdrawConstraint): rates are the exact-degree polynomial
    # through the pillars.
    POLYNOMIAL = 3


def fit_polynomial(x: tp.Sequence[float], y: tp.Sequence[float]) -> tp.Tuple[float, ...]:
    """Fit.Polynomial(x, y, len(x) - 1) (PolynomialInjectWithdrawConstraint.cs:66-72): least squares on the Vandermonde
    matrix — exact interpolation at this order.  Coefficients in ascending powers, like MathNet's Polynomial."""
    design = np.vander(np.asarray(x, dtype=np.float64), len(x), increasing=True)
    coef, *_ = np.linalg.lstsq(design, np.asarray(y, dtype=np.float64), rcond=None)
    return tuple(float(c) for c in coef)


def _horner(coef: tp.Sequence[float], z: float) -> float:   # Polynomial.Evaluate
    total = coef[-1]
    for c in reversed(coef[:-1]):
        total = total * z + c
    return total


class _Constraint(tp.NamedTuple):
    kind: int
    inventory: tp.Tuple[float, ...]
    min_rate: tp.Tuple[float, ...]     # polynomial kind: coefficients (ascending powers) instead of pillar rates
    max_rate: tp.Tuple[float, ...]

    def range(self, inventory: float) -> InjectWithdrawRange:
        if self.kind == CONSTRAINT_CONSTANT:
            return InjectWithdrawRange(self.min_rate[0], self.max_rate[0])
        if self.kind == CONSTRAINT_POLYNOMIAL:   # PolynomialInjectWithdrawConstraint.cs:80-85
            return InjectWithdrawRange(_horner(self.min_rate, inventory), _horner(self.max_rate, inventory))
        x = self.inventory
        if self.kind == CONSTRAINT_STEP:    # StepInjectWithdrawConstraint.cs:72-79
            if inventory < x[0] or inventory > x[-1]:
                raise ValueError("Value of inventory is outside of the interval [{}, {}].".format(x[0], x[-1]))
            idx = bisect.bisect_right(x, inventory) - 1
            return InjectWithdrawRange(self.min_rate[idx], self.max_rate[idx])
        # PiecewiseLinearInjectWithdrawConstraint.cs:67-72 (MathNet LinearSpline, left segment, linear extrapolation)
        k = min(max(bisect.bisect_right(x, inventory) - 1, 0), len(x) - 2)

        def interp(y):
            return y[k] + (inventory - x[k]) * ((y[k + 1] - y[k]) / (x[k + 1] - x[k]))
        return InjectWithdrawRange(interp(self.min_rate), interp(self.max_rate))


def _per_period(value, name):
    """scalar | Series | None -> callable(period) -> float"""
    if value is None:
        return lambda period: 0.0
    if utils.is_scalar(value):
        v = float(value)
        return lambda period: v
    if isinstance(value, pd.Series):
        def lookup(period):
            try:
                return float(value[period])
            except KeyError:
                raise ValueError("No {} provided for period {}.".format(name, period))
        return lookup
    raise TypeError("{} must be a number or pandas Series".format(name))


class CmdtyStorage:

    def __init__(self,
                 freq: str,
                 storage_start: utils.TimePeriodSpecType,
                 storage_end: utils.TimePeriodSpecType,
                 injection_cost: tp.Union[float, pd.Series],
                 withdrawal_cost: tp.Union[float, pd.Series],
                 ratchets=None,
                 ratchet_interp: tp.Optional[RatchetInterp] = None,
                 min_inventory: tp.Union[None, float, int, pd.Series] = None,
                 max_inventory: tp.Union[None, float, int, pd.Series] = None,
                 max_injection_rate: tp.Union[None, float, int, pd.Series] = None,
                 max_withdrawal_rate: tp.Union[None, float, int, pd.Series] = None,
                 cmdty_consumed_inject: tp.Union[None, float, int, pd.Series] = None,
                 cmdty_consumed_withdraw: tp.Union[None, float, int, pd.Series] = None,
                 terminal_storage_npv: tp.Optional[tp.Callable[[float, float], float]] = None,
                 inventory_loss: tp.Union[None, float, int, pd.Series] = None,
                 inventory_cost: tp.Union[None, float, int, pd.Series] = None):
        self._pfreq = utils.pandas_freq(freq)
        self._freq = freq
        self._start = utils.to_period(storage_start, freq)
        self._end = utils.to_period(storage_end, freq)
        if self._start >= self._end:    # CmdtyStorage.cs:196-197
            raise ValueError("Storage start period must be before end period.")

        self._ratchet_periods: tp.List[pd.Period] = []
        self._ratchet_constraints: tp.List[_Constraint] = []
        self._rate_lookup = None
        if ratchets is not None:
            utils.raise_if_not_none(min_inventory, "min_inventory parameter should not be provided if ratchets parameter is provided.")
            utils.raise_if_not_none(max_inventory, "max_inventory parameter should not be provided if ratchets parameter is provided.")
            utils.raise_if_not_none(max_injection_rate, "max_injection_rate parameter should not be provided if ratchets parameter is provided.")
            utils.raise_if_not_none(max_withdrawal_rate, "max_withdrawal_rate parameter should not be provided if ratchets parameter is provided.")
            utils.raise_if_none(ratchet_interp, "ratchet_interp parameter should be provided if ratchets parameter is provided.")
            self._init_ratchets(ratchets, ratchet_interp)
            if ratchet_interp == RatchetInterp.STEP and terminal_storage_npv is None:
                logger.warning("When ratchet_interp is RatchetInterp.STEP it is advisable to specify "
                               "terminal_storage_npv otherwise exceptions are likely to occur during valuation.")
        else:
            utils.raise_if_not_none(ratchet_interp, "ratchet_interp should not be provided if ratchets parameter is not provided.")
            utils.raise_if_none(min_inventory, "min_inventory parameter should be provided if ratchets parameter is not provided.")
            utils.raise_if_none(max_inventory, "max_inventory parameter should be provided if ratchets parameter is not provided.")
            utils.raise_if_none(max_injection_rate, "max_injection_rate parameter should be provided if ratchets parameter is not provided.")
            utils.raise_if_none(max_withdrawal_rate, "max_withdrawal_rate parameter should be provided if ratchets parameter is not provided.")
            if not utils.is_scalar(min_inventory) or min_inventory >= 0:
                pass
            if utils.is_scalar(min_inventory) and min_inventory < 0:
                raise ValueError("Minimum inventory must be non-negative.")
            if utils.is_scalar(max_inventory) and max_inventory < 0:
                raise ValueError("Maximum inventory must be non-negative.")
            inj = _per_period(max_injection_rate, "max_injection_rate")
            wdr = _per_period(max_withdrawal_rate, "max_withdrawal_rate")
            self._rate_lookup = lambda period: (-wdr(period), inj(period))
            self._min_inventory = _per_period(min_inventory, "min_inventory")
            self._max_inventory_raw = _per_period(max_inventory, "max_inventory")

        if utils.is_scalar(injection_cost) and injection_cost < 0:
            raise ValueError("Per unit inject cost must be non-negative.")
        if utils.is_scalar(withdrawal_cost) and withdrawal_cost < 0:
            raise ValueError("Per unit withdrawal cost must be non-negative.")
        self._injection_cost = _per_period(injection_cost, "injection_cost")
        self._withdrawal_cost = _per_period(withdrawal_cost, "withdrawal_cost")
        self._consumed_inject = _per_period(cmdty_consumed_inject, "cmdty_consumed_inject")
        self._consumed_withdraw = _per_period(cmdty_consumed_withdraw, "cmdty_consumed_withdraw")
        self._inventory_loss = _per_period(inventory_loss, "inventory_loss")
        self._inventory_cost = _per_period(inventory_cost, "inventory_cost")
        self._terminal_storage_npv = terminal_storage_npv
        self._empty_at_end = terminal_storage_npv is None

    # -- ratchets: CmdtyStorageBuilderExtensions.AddInjectWithdrawRanges (:142-257) -------------------------------
    def _init_ratchets(self, ratchets, ratchet_interp):
        entries = []
        for period, rates_by_inventory in ratchets:
            rows = [(float(i), float(lo), float(hi)) for i, lo, hi in rates_by_inventory]
            if len(rows) < 2:
                raise ValueError("Period {} contains less than 2 inject/withdraw/inventory constraints.".format(period))
            entries.append((utils.to_period(period, self._freq), rows))
        if not entries:
            raise ValueError("No inject/withdraw constrains provided.")
        entries.sort(key=lambda e: e[0])
        for (p0, _), (p1, _) in zip(entries, entries[1:]):
            if p0 == p1:
                raise ValueError("Repeated periods found in inject/withdraw ranges.")
        for period, rows in entries:
            min_inv = min(r[0] for r in rows)
            max_inv = max(r[0] for r in rows)
            if len(rows) == 2 and rows[0][1] == rows[1][1] and rows[0][2] == rows[1][2]:
                c = _Constraint(CONSTRAINT_CONSTANT, (min_inv, max_inv), (rows[0][1],) * 2, (rows[0][2],) * 2)   # :163-171
            else:
                rows = sorted(rows, key=lambda r: r[0])
                if ratchet_interp == RatchetInterp.POLYNOMIAL:
                    if len(rows) > 16:
                        raise ValueError("Polynomial ratchets support at most 16 inventory pillars.")
                    xs = [r[0] for r in rows]
                    c = _Constraint(CONSTRAINT_POLYNOMIAL, tuple(xs), fit_polynomial(xs, [r[1] for r in rows]),
                                    fit_polynomial(xs, [r[2] for r in rows]))
                else:
                    kind = CONSTRAINT_PIECEWISE_LINEAR if ratchet_interp == RatchetInterp.LINEAR else CONSTRAINT_STEP
                    if kind == CONSTRAINT_STEP:
                        self._validate_step(rows)
                    c = _Constraint(kind, tuple(r[0] for r in rows), tuple(r[1] for r in rows), tuple(r[2] for r in rows))
            self._ratchet_periods.append(period)
            self._ratchet_constraints.append(c)

    @staticmethod
    def _validate_step(rows):   # StepInjectWithdrawConstraint.cs:46-66
        if abs(rows[-2][2] - rows[-1][2]) > 1e-12:
            raise ValueError("Top two ratchets do not have he same max injection rate.")
        if abs(rows[-2][1] - rows[-1][1]) > 1e-12:
            raise ValueError("Top two ratchets do not have he same max withdrawal rate.")
        for i in range(1, len(rows) - 1):
            if rows[i][2] > rows[i - 1][2]:
                raise ValueError("Ratchet injection rates cannot increase with inventory.")
            if rows[i][1] > rows[i - 1][1]:
                raise ValueError("Ratchet withdrawal rates cannot decrease with inventory.")

    def _ratchet_index(self, period: pd.Period) -> int:
        idx = bisect.bisect_right(self._ratchet_periods, period) - 1
        if idx < 0:
            raise ValueError("No inject/withdraw constraint for period {} (before first ratchet period).".format(period))
        return idx

    def constraint_for(self, period: pd.Period) -> _Constraint:
        """_injectWithdrawConstraints(period) (CmdtyStorage.cs:99)."""
        if self._rate_lookup is not None:
            lo, hi = self._rate_lookup(period)
            return _Constraint(CONSTRAINT_CONSTANT, (0.0,), (lo,), (hi,))
        return self._ratchet_constraints[self._ratchet_index(period)]

    # -- mirrored public surface (cmdty_storage.py:200-277) -------------------------------------------------------
    def _period(self, period) -> pd.Period:
        return utils.to_period(period, self._freq)

    @property
    def freq(self) -> str:
        return self._freq

    @property
    def empty_at_end(self) -> bool:
        return self._empty_at_end

    @property
    def start(self) -> pd.Period:
        return self._start

    @property
    def end(self) -> pd.Period:
        return self._end

    def inject_withdraw_range(self, period, inventory) -> InjectWithdrawRange:
        period = self._period(period)
        mn, mx = self.min_inventory(period), self.max_inventory(period)
        if inventory < mn:    # CmdtyStorage.cs:88-94
            raise ValueError("Inventory of {} is below minimum allowed value of {} during period {}.".format(inventory, mn, period))
        if inventory > mx:
            raise ValueError("Inventory of {} above maximum allowed value of {} during period {}.".format(inventory, mx, period))
        if period >= self._end:
            return InjectWithdrawRange(0.0, 0.0)
        return self.constraint_for(period).range(inventory)

    def min_inventory(self, period) -> float:
        period = self._period(period)
        if self._rate_lookup is not None:
            return self._min_inventory(period)
        return self._ratchet_constraints[self._ratchet_index(period)].inventory[0]

    def max_inventory(self, period) -> float:
        period = self._period(period)
        if self._empty_at_end and period >= self._end:    # CmdtyStorage.cs:432-435
            return 0.0
        if self._rate_lookup is not None:
            return self._max_inventory_raw(period)
        return self._ratchet_constraints[self._ratchet_index(period)].inventory[-1]

    def injection_cost(self, period, inventory, injected_volume) -> float:
        return self._injection_cost(self._period(period)) * injected_volume

    def cmdty_consumed_inject(self, period, inventory, injected_volume) -> float:
        return self._consumed_inject(self._period(period)) * abs(injected_volume)

    def withdrawal_cost(self, period, inventory, withdrawn_volume) -> float:
        return self._withdrawal_cost(self._period(period)) * abs(withdrawn_volume)

    def cmdty_consumed_withdraw(self, period, inventory, withdrawn_volume) -> float:
        return self._consumed_withdraw(self._period(period)) * abs(withdrawn_volume)

    def terminal_storage_npv(self, cmdty_price, terminal_inventory) -> float:
        if self._terminal_storage_npv is None:
            return 0.0
        return self._terminal_storage_npv(cmdty_price, terminal_inventory)

    def inventory_pcnt_loss(self, period) -> float:
        return self._inventory_loss(self._period(period))

    def inventory_cost(self, period, inventory) -> float:
        return inventory * self._inventory_cost(self._period(period))

    # -- per-unit accessors used by the flattener ----------------------------------------------------------------
    def unit_injection_cost(self, period): return self._injection_cost(period)
    def unit_withdrawal_cost(self, period): return self._withdrawal_cost(period)
    def pcnt_consumed_inject(self, period): return self._consumed_inject(period)
    def pcnt_consumed_withdraw(self, period): return self._consumed_withdraw(period)
    def unit_inventory_cost(self, period): return self._inventory_cost(period)

    def terminal_coefficients(self) -> tp.Tuple[float, float, float, float]:
        """Bilinear form c0 + c1*price + c2*inv + c3*price*inv of terminal_storage_npv (device descriptor,
        SURVEY §7.3-3).  Verified by probing; anything else is unsupported on the device path."""
        if self._terminal_storage_npv is None:
            return 0.0, 0.0, 0.0, 0.0
        f = self._terminal_storage_npv
        c0 = float(f(0.0, 0.0))
        c1 = float(f(1.0, 0.0)) - c0
        c2 = float(f(0.0, 1.0)) - c0
        c3 = float(f(1.0, 1.0)) - c0 - c1 - c2
        for price, inv in ((37.25, 1234.5), (101.0, 0.75), (3.0, 52000.0)):
            expect = c0 + c1 * price + c2 * inv + c3 * price * inv
            got = float(f(price, inv))
            if abs(got - expect) > 1e-9 * max(1.0, abs(expect)):
                raise NotImplementedError("terminal_storage_npv is not of the form c0 + c1*price + c2*inventory + "
                                          "c3*price*inventory; no device descriptor exists for it (LSMC_EUNSUPPORTED).")
        return c0, c1, c2, c3
