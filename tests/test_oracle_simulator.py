"""Pins the oracle's multi-factor simulator.  The arithmetic of the reference's simulator lives in the un-vendored
NuGet package Cmdty.Core 0.1.0-alpha2020101802 (Cmdty.Storage.csproj:19), so bit-level parity is UNPINNED; what the
reference publishes in-tree is the model, and that is what these tests hold the oracle to:
  * one-factor log-variance  sigma^2 (1 - e^{-2 a t}) / (2 a)            (TestHelper.cs:77-86)
  * integrated covariances of log-forwards                               (multi_factor.py:166-187,225-229)
  * E[spot] = forward (martingale)                                        (LSMCTest:569-681 rely on it)
  * three-factor seasonal parameterisation                               (multi_factor.py:258-289)
  * Philox4x32-10 block function against the Random123 known-answer vectors (Salmon et al., SC'11, kat_vectors)."""
import math

import numpy as np
import pandas as pd
import pytest

from storage_b200._abi import Rng
from storage_b200.multi_factor_params import create_3_factor_season_params
from oracle import oracle
from fixtures import reference_problem, sin_forward_curve, STORAGE_END, VAL_DATE


def test_philox_known_answers():
    assert oracle.philox4x32_10([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert oracle.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert oracle.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_philox_normals_are_standard():
    z = np.array([oracle.philox_normal_pair(11, d, 3, 0) for d in range(20000)]).ravel()
    assert abs(z.mean()) < 4 / math.sqrt(z.size)
    assert abs(z.var() - 1.0) < 0.03
    assert abs(((z - z.mean()) ** 4).mean() - 3.0) < 0.15
    a = np.array([oracle.philox_normal_pair(11, d, 3, 0) for d in range(4000)])
    assert abs(np.corrcoef(a[:, 0], a[:, 1])[0, 1]) < 0.06


def test_one_factor_log_variance_and_martingale():
    N = 40000
    pb = reference_problem(num_sims=N)
    spot, fac = oracle.simulate(pb, Rng.philox(7, antithetic=False))
    s0 = int(pb.first_period_is_current)
    fwd = pb.forward[s0:]
    t = pb.sim_time[s0:]
    a, sigma = 12.5, 0.95
    for j in (0, 5, 60, len(t) - 1):
        ln = np.log(spot[j] / fwd[j])
        var = sigma ** 2 * (1 - math.exp(-2 * a * t[j])) / (2 * a)    # TestHelper.cs:82-84
        assert abs(ln.var() / var - 1) < 0.03
        se = math.sqrt((math.exp(var) - 1) / N)
        assert abs(spot[j].mean() / fwd[j] - 1) < 4 * se + 1e-12
        # Markov factor is the unit-vol OU state: Var = (1 - e^{-2at}) / 2a
        assert abs(fac[j, 0].var() / ((1 - math.exp(-2 * a * t[j])) / (2 * a)) - 1) < 0.03


def test_antithetic_pairs_are_adjacent_and_mirror():
    pb = reference_problem(num_sims=64)
    spot, fac = oracle.simulate(pb, Rng.philox(3, antithetic=True))
    assert np.array_equal(fac[:, :, 0::2], -fac[:, :, 1::2])
    s0 = int(pb.first_period_is_current)
    # spot = F exp(v x - var/2): mirrored pair multiplies to (F e^{-var/2})^2
    prod = spot[:, 0::2] * spot[:, 1::2]
    assert np.allclose(prod, prod[:, :1], rtol=1e-12)


def test_sharded_philox_equals_global():
    """Path n of a shard that starts at sim_offset draws what global path sim_offset + n draws (multi-GPU sharding)."""
    pb = reference_problem(num_sims=96)
    full, full_f = oracle.simulate(pb, Rng.philox(21, antithetic=True))
    part = pb.copy(n_sims=32, sim_offset=64, n_sims_global=96)
    sh, sh_f = oracle.simulate(part, Rng.philox(21, antithetic=True))
    assert np.array_equal(sh, full[:, 64:]) and np.array_equal(sh_f, full_f[:, :, 64:])


def test_multi_factor_covariance_matches_published_formula():
    """multi_factor.py:166-187: integrated covariance of log forwards; here for spot (T1 = T2 = t) across two dates."""
    idx = pd.period_range(VAL_DATE, STORAGE_END, freq="D")
    mrs, vols, rho = [0.0, 16.2, 45.0], [0.25, 0.64, 0.45], 0.35
    factors = [(m, pd.Series(v, index=idx)) for m, v in zip(mrs, vols)]
    corr = np.array([[rho ** abs(i - j) for j in range(3)] for i in range(3)])
    N = 60000
    pb = reference_problem(num_sims=N, factors=factors, corrs=corr, basis="1 + x0 + x1 + x2")
    spot, fac = oracle.simulate(pb, Rng.philox(5, antithetic=False))
    s0 = int(pb.first_period_is_current)
    t = pb.sim_time[s0:]
    fwd = pb.forward[s0:]

    def cont_ext(c, dt):
        return dt if c == 0 else (1 - math.exp(-c * dt)) / c

    for j1, j2 in ((10, 10), (10, 40), (100, 180)):
        t1, t2 = t[j1], t[j2]
        cov = sum(vols[i] * vols[k] * corr[i, k] * math.exp(-mrs[i] * (t1 - t1) - mrs[k] * (t2 - t1)) * cont_ext(mrs[i] + mrs[k], t1)
                  for i in range(3) for k in range(3))
        l1, l2 = np.log(spot[j1] / fwd[j1]), np.log(spot[j2] / fwd[j2])
        got = np.cov(l1, l2)[0, 1]
        v1 = sum(vols[i] * vols[k] * corr[i, k] * cont_ext(mrs[i] + mrs[k], t1) for i in range(3) for k in range(3))
        v2 = sum(vols[i] * vols[k] * corr[i, k] * cont_ext(mrs[i] + mrs[k], t2) for i in range(3) for k in range(3))
        assert abs(got - cov) < 5 * math.sqrt((v1 * v2 + cov * cov) / N)
    # factor cross-covariance at one date: rho_ik (1 - e^{-(a_i+a_k) t}) / (a_i + a_k)
    j = 120
    for i in range(3):
        for k in range(i, 3):
            expect = corr[i, k] * cont_ext(mrs[i] + mrs[k], t[j])
            got = np.cov(fac[j, i], fac[j, k])[0, 1]
            vi, vk = cont_ext(2 * mrs[i], t[j]), cont_ext(2 * mrs[k], t[j])
            assert abs(got - expect) < 5 * math.sqrt((vi * vk + expect * expect) / N)


def test_three_factor_seasonal_parameters():   # multi_factor.py:258-289
    factors, corr = create_3_factor_season_params("D", 91.0, 0.85, 0.30, 0.19, "2021-04-01", "2022-04-01")
    assert np.array_equal(corr, np.eye(3))
    assert [f[0] for f in factors] == [91.0, 0.0, 0.0]
    assert set(factors[0][1].values) == {0.85} and set(factors[1][1].values) == {0.30}
    seasonal = factors[2][1]
    peak = pd.Period("2021-02-01", freq="D")
    for day in ("2021-04-01", "2021-08-02", "2022-02-01"):
        per = pd.Period(day, freq="D")
        tt = (per.start_time - peak.start_time).total_seconds() / (365.25 * 86400)
        assert abs(seasonal[per] - 0.095 * math.sin(2 * math.pi * tt + math.pi / 2)) < 1e-15
    assert abs(seasonal[pd.Period("2022-02-01", freq="D")] - 0.095) < 1e-4   # one 365-day year after the peak


def test_injected_normals_reproduce_ou_recursion():
    """x(t+dt) = e^{-a dt} x(t) + L z with L = chol of the exact step covariance; spot = F exp(sum v x - var/2)."""
    pb = reference_problem(num_sims=4, mean_reversion=3.0, spot_vol=0.5)
    S = pb.n_sim_periods
    z = np.random.default_rng(1).standard_normal((S, 1, 4))
    spot, fac = oracle.simulate(pb, Rng.from_normals(z))
    s0 = int(pb.first_period_is_current)
    t = np.concatenate([[0.0], pb.sim_time[s0:]])
    x = np.zeros(4)
    for j in range(S):
        dt = t[j + 1] - t[j]
        x = math.exp(-3.0 * dt) * x + math.sqrt((1 - math.exp(-6.0 * dt)) / 6.0) * z[j, 0]
        np.testing.assert_allclose(fac[j, 0], x, rtol=1e-12, atol=1e-15)
        var = 0.25 * (1 - math.exp(-6.0 * t[j + 1])) / 6.0
        np.testing.assert_allclose(spot[j], pb.forward[s0 + j] * np.exp(0.5 * x - 0.5 * var), rtol=1e-12)


def test_multi_factor_model_closed_forms_match_the_simulator():
    """storage_b200.MultiFactorModel (the reference's analytic covariances, multi_factor.py:166-229) against the oracle's
    simulated log-spot: one-factor variance sigma^2 (1 - exp(-2 alpha T)) / (2 alpha) (TestHelper.cs:77-86) in closed form,
    two-factor variance and correlation against a 40 000-path sample."""
    import math
    import pandas as pd
    from storage_b200 import MultiFactorModel
    idx = pd.period_range("2019-08-29", "2020-04-01", freq="D")
    one = MultiFactorModel("D", [(12.5, pd.Series(0.95, index=idx))])
    expiry = pd.Period("2019-12-15", freq="D")
    T = (expiry.ordinal - idx[0].ordinal) / 365.0
    assert one.integrated_variance(idx[0], expiry, expiry) == pytest.approx(0.95 ** 2 * (1 - math.exp(-25.0 * T)) / 25.0, rel=1e-12)
    assert one.integrated_vol(idx[0], expiry, expiry) == pytest.approx(math.sqrt(0.95 ** 2 * (1 - math.exp(-25.0 * T)) / 25.0 / T), rel=1e-12)
    assert one.integrated_corr(idx[0], expiry, expiry, expiry) == 1.0
    two = MultiFactorModel("D", [(0.0, pd.Series(0.3, index=idx)), (16.0, pd.Series(1.1, index=idx))], 0.4)
    pb = reference_problem(num_sims=40_000, factors=[(0.0, pd.Series(0.3, index=idx)), (16.0, pd.Series(1.1, index=idx))],
                           corrs=np.array([[1.0, 0.4], [0.4, 1.0]]), basis="1 + x0 + x1")
    spot, _ = oracle.simulate(pb, Rng.philox(5))
    for j in (3, 60, 200):
        period = pb.periods[j + (pb.n_periods - pb.n_sim_periods)]
        sample = np.log(spot[j]).var()
        assert sample == pytest.approx(two.integrated_variance(idx[0], period, period), rel=0.03)
