"""MultiFactorModel — closed-form covariances of log forward prices under the multi-factor model the LSMC path simulates
(same public methods as the reference's cmdty_storage.MultiFactorModel, multi_factor.py:149-250):

    dF(t,T)/F(t,T) = sum_i sigma_i(T) exp(-alpha_i (T - t)) dz_i(t),   corr(dz_i, dz_j) = rho_ij

so over an observation window [0, tau] (times measured from obs_start by `time_func`, Act/365 by default)

    Cov[ln F(., T1), ln F(., T2)] = sum_ij sigma_i(T1) sigma_j(T2) rho_ij exp(-alpha_i t1 - alpha_j t2) E(alpha_i + alpha_j, tau)
    E(a, tau) = (exp(a tau) - 1) / a        (tau when a = 0)

It is the analytic counterpart of the K1 simulation kernel and is what tests/test_oracle_simulator.py holds the simulator
to; nothing on the device path calls it."""
import math
import typing as tp

import numpy as np

from . import utils
from .multi_factor_params import create_3_factor_season_params

TimeFunctionType = tp.Callable[[tp.Any, tp.Any], float]


def _growth_integral(rate_sum: float, tau: float) -> float:
    return tau if rate_sum == 0.0 else math.expm1(rate_sum * tau) / rate_sum


class MultiFactorModel:
    _corr_tolerance = 1E-10

    def __init__(self, freq: str, factors, factor_corrs=None, time_func: tp.Optional[TimeFunctionType] = None):
        from .multi_factor import _validate_multi_factor_params
        self._factors = list(factors)
        self._corrs = _validate_multi_factor_params(self._factors, factor_corrs)
        self._time = utils.act_365 if time_func is None else time_func
        self._freq = freq
        for k, (mean_reversion, _) in enumerate(self._factors):
            if mean_reversion < 0.0:
                raise ValueError("Mean reversion value of {} for factor at index {} not valid as is negative.".format(mean_reversion, k))

    def _vol(self, k: int, contract) -> float:
        curve = self._factors[k][1]
        vol = curve.get(contract, None)
        if vol is None:
            raise ValueError("No point in vol curve of factor {} for fwd_contract_1 value of {}.".format(k, contract))
        return float(vol)

    def integrated_covar(self, obs_start, obs_end, fwd_contract_1, fwd_contract_2) -> float:
        tau = self._time(obs_start, obs_end)
        if tau < 0.0:
            raise ValueError("obs_end cannot be before obs_start.")
        t1, t2 = self._time(obs_start, fwd_contract_1), self._time(obs_start, fwd_contract_2)
        n = len(self._factors)
        alpha = np.array([f[0] for f in self._factors], dtype=np.float64)
        load_1 = np.array([self._vol(k, fwd_contract_1) for k in range(n)]) * np.exp(-alpha * t1)
        load_2 = np.array([self._vol(k, fwd_contract_2) for k in range(n)]) * np.exp(-alpha * t2)
        growth = np.array([[_growth_integral(alpha[i] + alpha[j], tau) for j in range(n)] for i in range(n)])
        return float(load_1 @ (self._corrs * growth) @ load_2)

    def integrated_variance(self, obs_start, obs_end, fwd_contract) -> float:
        return self.integrated_covar(obs_start, obs_end, fwd_contract, fwd_contract)

    def integrated_stan_dev(self, obs_start, obs_end, fwd_contract) -> float:
        return math.sqrt(self.integrated_variance(obs_start, obs_end, fwd_contract))

    def integrated_vol(self, val_date, expiry, fwd_contract) -> float:
        time_to_expiry = self._time(val_date, expiry)
        if time_to_expiry <= 0:
            raise ValueError("val_date must be before expiry.")
        return math.sqrt(self.integrated_variance(val_date, expiry, fwd_contract) / time_to_expiry)

    def integrated_corr(self, obs_start, obs_end, fwd_contract_1, fwd_contract_2) -> float:
        cov = self.integrated_covar(obs_start, obs_end, fwd_contract_1, fwd_contract_2)
        corr = cov / math.sqrt(self.integrated_variance(obs_start, obs_end, fwd_contract_1) *
                               self.integrated_variance(obs_start, obs_end, fwd_contract_2))
        if 1.0 < corr < 1.0 + self._corr_tolerance:
            return 1.0
        if -1.0 - self._corr_tolerance < corr < -1.0:
            return -1.0
        return corr

    @staticmethod
    def for_3_factor_seasonal(freq: str, spot_mean_reversion: float, spot_vol: float, long_term_vol: float, seasonal_vol: float,
                              start, end, time_func: tp.Optional[TimeFunctionType] = None) -> 'MultiFactorModel':
        factors, corrs = create_3_factor_season_params(freq, spot_mean_reversion, spot_vol, long_term_vol, seasonal_vol, start, end)
        return MultiFactorModel(freq, factors, corrs, time_func)
