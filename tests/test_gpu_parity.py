"""Parity of the CUDA path (through the C ABI) with the CPU oracle.  Needs a B200: `pytest -m gpu`.

Bars (BASELINE.json north_star):
  * injected normals: simulated panels bit-exact; with the oracle's regression coefficients injected the whole value
    surface and every forward decision index are bit-exact; end to end (own normal-equation regression vs the oracle's
    Householder QR) NPV / deltas / profile / trigger prices within 1e-9 relative and no decision flips except between
    decisions of numerically identical volume;
  * device RNG: Philox stream identical to the oracle's, NPV within 3 standard errors.
"""
import json
import os

import numpy as np
import pytest

from storage_b200 import CmdtyStorage, RatchetInterp, _abi, _native
from storage_b200._abi import Rng
from storage_b200.configs import build_config, full_polynomial_basis
from oracle import oracle
from fixtures import (reference_problem, simple_daily_storage, ratchet_daily_storage, three_calls_problem,
                      forced_inject_withdraw_problem, STORAGE_START, STORAGE_END)

pytestmark = pytest.mark.gpu
REL = 1e-9   # north_star tolerance with injected normals


def anti_normals(seed, S, F, N):
    rs = np.random.default_rng(seed)
    z = rs.standard_normal((S, F, (N + 1) // 2))
    out = np.empty((S, F, N))
    out[:, :, 0::2] = z[:, :, :(N + 1) // 2]
    out[:, :, 1::2] = -z[:, :, :N // 2]
    return out


def rel_err(a, b, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(b)
    if not m.any():
        return 0.0
    s = scale if scale is not None else max(np.max(np.abs(b[m])), 1e-300)
    return float(np.max(np.abs(a[m] - b[m])) / s)


def material_flips(got, ref):
    """Decision differences that change the decided volume by more than 1e-9 of the largest rate."""
    scale = np.max(np.abs(ref.inject_withdraw))
    return int(np.sum(np.abs(got.inject_withdraw - ref.inject_withdraw) > 1e-9 * max(scale, 1e-300)))


def comparable_trigger_rows(got, ref, side):
    """Rows whose trigger side is well defined.  The reference emits a side only when `max > alternative`
    (LSV.cs:510,538); near the end of a must-be-empty storage both are the same forced volume up to rounding of the
    MEAN inventory (a sum over sims, whose order differs on the device), so presence itself is rounding noise there.
    Such degenerate rows (volume span below 1e-9 of the rate scale) are excluded and must stay rare."""
    gp, rp = (got.trigger_inject, ref.trigger_inject) if side == 0 else (got.trigger_withdraw, ref.trigger_withdraw)
    scale = max(np.nanmax(np.abs(rp[..., 0])) if np.any(~np.isnan(rp[..., 0])) else 0.0, 1e-300)

    def degenerate(prof):
        v = prof[..., 0]
        has = ~np.isnan(v[:, 0])
        span = np.where(has, np.nanmax(np.where(np.isnan(v), -np.inf, v), axis=1) - np.nanmin(np.where(np.isnan(v), np.inf, v), axis=1), np.inf)
        return has & (span < 1e-9 * scale)

    mismatch = np.isnan(gp[:, 0, 0]) != np.isnan(rp[:, 0, 0])
    excluded = mismatch & (degenerate(gp) | degenerate(rp))
    assert not np.any(mismatch & ~excluded), "trigger presence differs on a well-defined row"
    assert excluded.sum() <= max(2, len(excluded) // 50)
    return ~excluded


def step_ratchet_storage():
    """The reference's ratchet fixture (LSMCTest:91-125) under RatchetInterp.STEP (StepInjectWithdrawConstraint.cs:46-79:
    the top two pillars must carry the same rates; a terminal value is advised, cmdty_storage.py:149-152)."""
    ratchets = [
        (STORAGE_START, [(0.0, -702.7, 650.0), (15_000, -785.0, 552.5), (30_000, -790.6, 512.8), (40_000, -825.6, 498.6),
                         (52_500, -825.6, 498.6)]),
        ("2020-02-01", [(0.0, -645.35, 650.0), (13_000, -656.0, 552.5), (28_000, -689.6, 512.8), (42_000, -701.06, 498.6),
                        (52_500, -701.06, 498.6)]),
    ]
    return CmdtyStorage("D", STORAGE_START, STORAGE_END, 1.25, 0.93, ratchets=ratchets, ratchet_interp=RatchetInterp.STEP,
                        terminal_storage_npv=lambda price, inventory: 0.0)


def polynomial_ratchet_storage():
    """Rates that are polynomials in the inventory (PolynomialInjectWithdrawConstraint.cs:66-85; C# builder only): cubic in the
    first table, quadratic in the second; the backward inventory-space sweep needs the bound solver near the end (:87-156)."""
    ratchets = [
        (STORAGE_START, [(0.0, -30.0, 25.0), (500.0, -36.0, 21.0), (1000.0, -38.0, 19.5), (1400.0, -39.0, 19.0)]),
        ("2020-02-01", [(0.0, -28.0, 26.0), (700.0, -33.0, 22.0), (1400.0, -35.0, 20.5)]),
    ]
    return CmdtyStorage("D", STORAGE_START, STORAGE_END, 1.25, 0.93, ratchets=ratchets, ratchet_interp=RatchetInterp.POLYNOMIAL)


def inventory_cost_storage():
    """Non-zero per-unit inventory cost (CmdtyStorage.cs:544-567, LSV.cs:241-242): the kInvc variants of the sweeps."""
    return CmdtyStorage("D", STORAGE_START, STORAGE_END, 1.25, 0.93, min_inventory=0.0, max_inventory=52_500.0,
                        max_injection_rate=625.0, max_withdrawal_rate=850.0, inventory_cost=0.004)


# five factors, K = 16, full grid: the pipelined sweep's shared memory only fits with 40-path tiles (sweep4.cu sweep4_tile)
FIVE_FACTOR_K16 = ("1 + x0 + x1 + x2 + x3 + x4 + x0**2 + x1**2 + x2**2 + x3**2 + x4**2 + x0*x1 + x1*x2 + x2*x3 + x3*x4 + s")

CASES = {
    "c4_five_factor_k56": lambda: build_config("c4", num_sims=512),                              # K > 32: sweep_kernel MT=8, solve_kernel<64>, forward_kernel<64>
    "c3_cubic_basis_k20": lambda: build_config("c3", num_sims=640, basis=full_polynomial_basis(3, 3)),   # 16 < K <= 32: MT=4
    # grids beyond 104 rows: the general sweep for every period, two rounds of decision items per tile (and, with K > 16, the
    # basis program / x~ not overlapped: the three warps it would run on own decision items)
    "c3_fine_grid_130": lambda: build_config("c3", num_sims=512, end="2021-08-01", grid_points=130),
    "c3_cubic_basis_k20_grid_120": lambda: build_config("c3", num_sims=512, end="2021-07-01", basis=full_polynomial_basis(3, 3), grid_points=120),
    "c4_five_factor_k16_tp40": lambda: build_config("c4", num_sims=600, end="2021-07-01", basis=FIVE_FACTOR_K16),
    "reference_fixture_step_ratchets": lambda: reference_problem(storage=step_ratchet_storage(), num_sims=768, numerical_tolerance=1e-9),
    "inventory_cost": lambda: reference_problem(storage=inventory_cost_storage(), num_sims=512),
    "polynomial_ratchets": lambda: reference_problem(storage=polynomial_ratchet_storage(), num_sims=640, inventory=300.0, numerical_tolerance=1e-9),
    "c1_readme_simple_3f": lambda: build_config("c1", num_sims=1000),
    "c2_one_factor": lambda: build_config("c2", num_sims=1536),
    "c3_ratchets_losses_3f": lambda: build_config("c3", num_sims=1024),
    "c3_ragged_paths": lambda: build_config("c3", num_sims=333),
    "c5_battery_half_hourly": lambda: build_config("c5", num_sims=512),
    "reference_fixture_1f": lambda: reference_problem(num_sims=2000),
    "reference_fixture_ratchets": lambda: reference_problem(storage=ratchet_daily_storage(), num_sims=768, numerical_tolerance=1e-9),
    "terminal_value_not_empty": lambda: reference_problem(storage=simple_daily_storage(lambda p, i: p * i - 999.0), num_sims=512),
    "before_storage_start": lambda: reference_problem(val_date="2019-07-20", inventory=0.0, num_sims=512),
    "extra_decisions": lambda: reference_problem(num_sims=512, extra_decisions=2),
    "three_calls": lambda: three_calls_problem(num_sims=1024)[0],
    "forced_inject_withdraw": lambda: forced_inject_withdraw_problem(num_sims=1000)[0],   # LSMCTest:686-837: one-point grids
}


@pytest.fixture(scope="module")
def ctx():
    with _native.Context(0) as c:
        yield c


@pytest.mark.parametrize("case", sorted(CASES))
def test_staged_bit_exact_with_injected_normals_and_coefficients(case):
    pb = CASES[case]()
    S, F, P, K, N = pb.n_sim_periods, pb.n_factors, pb.n_periods, pb.n_basis, int(pb.n_sims)
    zr, zv = anti_normals(1, S, F, N), anti_normals(2, S, F, N)
    o = oracle.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), keep_values=True, keep_factors=True,
                         panels=("spot_regress", "spot_valuation", "inventory", "inject_withdraw", "pv"))
    with _native.Context(0) as ctx:
        ctx.request_outputs(panels=True, decision_index=True)
        ctx.debug_keep_values(True)
        ctx.set_problem(pb)
        lo, hi = ctx.inventory_space()
        assert np.array_equal(lo, o.inventory_space_min) and np.array_equal(hi, o.inventory_space_max)
        for p in range(P):
            assert np.array_equal(ctx.grid(p), o.grids[p, :o.grid_count[p]])
        ctx.simulate(0, Rng.from_normals(zr))
        assert np.array_equal(ctx.sim_panel(0), o.results.spot_regress)
        for f in range(F):
            assert np.array_equal(ctx.sim_panel(0, f), o.factors_regress[:, f, :])
        for p in range(P - 1):
            g = o.grid_count[p + 1]
            if p == 0 and pb.first_period_is_current:
                coeffs = np.zeros((g, K)); coeffs[:, 0] = o.current_continuation[:g]
            else:
                coeffs = o.regression_coeffs[p, :g, :]
            ctx.debug_set_regression_coeffs(p, coeffs)
        ctx.backward()
        for p in range(P - 1, -1, -1):
            assert np.array_equal(ctx.debug_values(p), o.values[p, :o.grid_count[p], :]), "value surface of period %d" % p
        ctx.simulate(1, Rng.from_normals(zv))
        ctx.forward()
        r = ctx.get_results(panels=("inventory", "inject_withdraw", "pv", "spot_valuation"), decision_index=True)
        assert np.array_equal(r.spot_valuation, o.results.spot_valuation)
        assert np.array_equal(r.decision_index, o.results.decision_index)        # bit-exact decisions
        assert np.array_equal(r.inject_withdraw, o.results.inject_withdraw)
        assert np.array_equal(r.inventory, o.results.inventory)
        assert np.array_equal(r.pv, o.results.pv)
        assert rel_err(r.npv, o.results.npv) < 1e-13                              # only the summation order differs
        assert rel_err(r.deltas, o.results.deltas) < 1e-12


# Trigger prices (LSV.cs:619-638) are differences of MEAN fitted continuation values divided by df * (volume change).  A mean
# fitted value is sum_k xbar_k beta_lk; where the basis is nearly collinear (the first half-hourly steps of C5: cond(X) ~ 1e14,
# sum_k |xbar_k beta_lk| = 4e10 against a value of 2e2) that sum cancels catastrophically, so the reference's own formula is
# only defined to eps * sum_k |xbar_k beta_lk| — any other summation order (MathNet's, the oracle's, the device's) moves the
# price by that much.  Every trigger price is therefore held to 1e-9 of the price scale OR to 16 roundings of that
# cancellation sum (divided by df * volume increment), whichever is larger; the metrics (profiles/r02_parity_metrics.json)
# record per case the largest error, the largest error over the well-conditioned periods, and how many periods exceed 1e-9.
METRICS = {}
EPS = 2.220446049250313e-16


@pytest.fixture(scope="module", autouse=True)
def dump_metrics():
    yield
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if METRICS and os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "parity_metrics.json"), "w") as f:
            json.dump(METRICS, f, indent=1, sort_keys=True)


def trigger_price_bounds(pb, o, price_scale):
    """Per decision period: the absolute tolerance of a trigger price (see above) and the conditioning figures behind it.
    Period p also inherits the noise of period p+1: its continuation values are regressions (or, in the current period,
    plain means) of V_{p+1}, whose backward decisions were taken on the fitted values of regression p+1 — where those are
    noise, near-ties resolve differently and move V_{p+1} by a realised, not a near-tied, difference."""
    P, S = pb.n_periods, pb.n_sim_periods
    noise = np.zeros(P)
    amp_max = 0.0
    for p in range(P - 1):
        j = p - (P - S)
        g = o.grid_count[p + 1]
        if j < 0:   # current period: the continuation is a stored mean, no regression
            continue
        X = oracle.design_matrix(pb, o.results.spot_valuation[j], o.factors_valuation[j])
        amp = float(np.max(np.abs(o.regression_coeffs[p, :g, :]) @ np.abs(X.mean(axis=0))))
        vols = np.concatenate([o.results.trigger_inject[p, :, 0], o.results.trigger_withdraw[p, :, 0]])
        vols = np.sort(vols[~np.isnan(vols)])
        steps = np.diff(vols)
        steps = steps[steps > 0]
        inc = float(np.min(steps)) if len(steps) else float(pb.grid_spacing)
        noise[p] = 16 * EPS * amp / (pb.settle_df[p] * inc)
        amp_max = max(amp_max, amp)
    tol = np.maximum(REL * price_scale, np.maximum(noise[:-1], noise[1:]))
    return tol, amp_max, int(np.sum(tol > REL * price_scale))


@pytest.mark.parametrize("case", sorted(CASES))
def test_end_to_end_within_1e9_with_injected_normals(case, ctx):
    pb = CASES[case]()
    S, F, N = pb.n_sim_periods, pb.n_factors, int(pb.n_sims)
    zr, zv = anti_normals(3, S, F, N), anti_normals(4, S, F, N)
    o = oracle.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), keep_factors=True, panels=("inject_withdraw", "spot_valuation"))
    ref = o.results
    got = ctx.calculate(pb, Rng.from_normals(zr), Rng.from_normals(zv), panels=("inject_withdraw",), decision_index=True)
    price_scale = float(np.max(pb.forward))
    price_tol, amp_max, noisy = trigger_price_bounds(pb, o, price_scale)
    m = {"npv_rel": rel_err(got.npv, ref.npv), "backward_npv_rel": rel_err(got.backward_npv, ref.backward_npv),
         "delta_rel": rel_err(got.deltas, ref.deltas), "profile_rel": rel_err(got.profile, ref.profile),
         "decision_flips": material_flips(got, ref),
         "index_differences": int(np.sum(got.decision_index != ref.decision_index)),
         "pv_by_sim_rel": rel_err(got.pv_by_sim, ref.pv_by_sim, max(abs(ref.npv), 1.0)),
         "trigger_volume_rel": 0.0, "trigger_price_rel": 0.0, "trigger_price_rel_well_conditioned": 0.0,
         "trigger_periods_noise_limited": noisy, "max_sum_abs_xbar_beta": amp_max,
         "paths": N, "periods": pb.n_periods - 1, "basis": pb.n_basis}
    METRICS[case] = m
    price_ok = True
    for side, (gp, rp) in enumerate(((got.trigger_inject, ref.trigger_inject), (got.trigger_withdraw, ref.trigger_withdraw))):
        rows = comparable_trigger_rows(got, ref, side)
        cols = [2 * side, 2 * side + 1]
        m["trigger_volume_rel"] = max(m["trigger_volume_rel"], rel_err(got.trigger_summary[rows][:, cols[0]], ref.trigger_summary[rows][:, cols[0]]),
                                      rel_err(gp[rows][..., 0], rp[rows][..., 0]))
        m["trigger_rows_excluded"] = m.get("trigger_rows_excluded", 0) + int((~rows).sum())
        for prices_got, prices_ref in ((got.trigger_summary[:, cols[1]][:, None], ref.trigger_summary[:, cols[1]][:, None]), (gp[..., 1], rp[..., 1])):
            err = np.abs(prices_got - prices_ref)
            err = np.where(np.isnan(prices_ref) | ~rows[:, None], 0.0, err)
            assert not np.any(np.isnan(err))
            per_period = err.max(axis=1)
            m["trigger_price_rel"] = max(m["trigger_price_rel"], float(per_period.max()) / price_scale)
            tight = price_tol <= REL * price_scale
            if tight.any():
                m["trigger_price_rel_well_conditioned"] = max(m["trigger_price_rel_well_conditioned"], float(per_period[tight].max()) / price_scale)
            price_ok = price_ok and bool(np.all(per_period <= price_tol))
            m["trigger_periods_above_1e9"] = max(m.get("trigger_periods_above_1e9", 0), int(np.sum(per_period > REL * price_scale)))
    assert m["npv_rel"] < REL and m["backward_npv_rel"] < REL
    assert m["delta_rel"] < REL and m["profile_rel"] < REL
    assert m["decision_flips"] == 0
    assert m["trigger_volume_rel"] < REL
    assert price_ok, "a trigger price is outside 1e-9 of the price scale and outside the cancellation noise of its period"
    assert m["trigger_price_rel_well_conditioned"] < REL
    assert m["pv_by_sim_rel"] < REL


@pytest.mark.parametrize("case", ["c2_one_factor", "c3_ratchets_losses_3f", "c3_ragged_paths", "reference_fixture_1f", "c3_cubic_basis_k20"])
def test_device_rng_matches_oracle_stream_and_npv_within_3_standard_errors(case, ctx):
    pb = CASES[case]()
    ref = oracle.calculate(pb, Rng.philox(11), Rng.philox(22), panels=("spot_regress", "spot_valuation")).results
    got = ctx.calculate(pb, Rng.philox(11), Rng.philox(22), panels=("spot_regress", "spot_valuation"))
    # same Philox counters and Box-Muller; log/sincos come from different math libraries -> a few ulps
    assert rel_err(got.spot_regress, ref.spot_regress) < 1e-12
    assert rel_err(got.spot_valuation, ref.spot_valuation) < 1e-12
    assert abs(got.npv - ref.npv) < 3 * ref.npv_std_error
    assert rel_err(got.npv_std_error, ref.npv_std_error) < 1e-3
    # independent draws on the device: still within 3 standard errors (of the difference)
    other = ctx.calculate(pb, Rng.philox(101), Rng.philox(202))
    assert abs(other.npv - ref.npv) < 3 * np.hypot(ref.npv_std_error, other.npv_std_error) + 1e-9


def test_panels_mode_accepts_external_simulations(ctx):
    """LSMC_RNG_PANELS: spot/factor panels from another simulator (e.g. Cmdty.Core's own) are used as they are."""
    pb = CASES["reference_fixture_1f"]().copy(n_sims=512)
    spot_r, fac_r = oracle.simulate(pb, Rng.philox(5))
    spot_v, fac_v = oracle.simulate(pb, Rng.philox(6))
    ref = oracle.calculate(pb, Rng.from_panels(spot_r, fac_r), Rng.from_panels(spot_v, fac_v)).results
    got = ctx.calculate(pb, Rng.from_panels(spot_r, fac_r), Rng.from_panels(spot_v, fac_v))
    assert rel_err(got.npv, ref.npv) < REL and rel_err(got.deltas, ref.deltas) < REL


def test_full_size_properties_c3(ctx):
    """Size-independent properties at BASELINE's full C3 size (1M paths) where the oracle cannot follow."""
    pb = build_config("c3", num_sims=1_000_000)
    res = ctx.calculate(pb, Rng.philox(11), Rng.philox(22))
    small = oracle.calculate(build_config("c3", num_sims=16384), Rng.philox(11), Rng.philox(22)).results
    assert abs(res.npv - small.npv) < 4 * np.hypot(res.npv_std_error, small.npv_std_error)
    assert abs(res.backward_npv - res.npv) < 0.01 * abs(res.npv)           # in-sample vs out-of-sample value
    prof = res.profile
    assert prof[0, 0] == pb.start_inventory and abs(prof[-1, 0]) < 1e-6    # must be empty at end
    # expected inventory recursion: E[v_{t+1}] = E[v_t] + E[d_t] - E[loss_t]
    assert np.allclose(prof[1:, 0], prof[:-1, 0] + prof[:-1, 1] - prof[:-1, 3], rtol=0, atol=1e-6)
    assert np.allclose(prof[:, 4], -prof[:, 1] - prof[:, 2], rtol=0, atol=1e-9)   # net volume identity
    assert abs(prof[:, 5].sum() - res.npv) < 1e-6 * abs(res.npv)                 # NPV = sum of expected period PVs
    assert np.all(prof[:, 0] >= -1e-9) and np.all(prof[:, 0] <= pb.max_inventory.max() + 1e-9)
    again = ctx.calculate(pb, Rng.philox(11), Rng.philox(22))
    assert again.npv == res.npv and np.array_equal(again.deltas, res.deltas)      # deterministic reductions


@pytest.mark.parametrize("config,paths", [("c3", 150_000), ("c2", 100_000), ("c5", 20_000), ("c1", 50_000), ("c3", 1001)])
def test_forward_streaming_and_straight_line_kernels_equal_the_bisecting_kernel(config, paths, monkeypatch):
    """forward2_kernel (streamed per-period constants, O(1) grid lookup from the class table that replays SH.cs:280-314) and
    its predecessor forward_fast_kernel against forward_kernel's bisection on every path of a large run: decisions,
    inventories and PVs must be bit-identical."""
    pb = build_config(config, num_sims=paths)
    out = {}
    for mode in ("2", "1", "0"):   # LSMC_FORWARD (read in lsmc_set_problem): 2 streaming, 1 straight-line, 0 bisecting kernel
        monkeypatch.setenv("LSMC_FORWARD", mode)
        with _native.Context(0) as c:
            out[mode] = c.calculate(pb, Rng.philox(11), Rng.philox(22), panels=("inventory", "inject_withdraw", "pv"), decision_index=True)
    slow = out["0"]
    for mode in ("2", "1"):
        fast = out[mode]
        assert np.array_equal(fast.decision_index, slow.decision_index), mode
        assert np.array_equal(fast.inventory, slow.inventory), mode
        assert np.array_equal(fast.inject_withdraw, slow.inject_withdraw), mode
        assert np.array_equal(fast.pv, slow.pv), mode
        assert rel_err(fast.npv, slow.npv) < 1e-13   # partition of the sums differs
        assert rel_err(fast.deltas, slow.deltas) < 1e-12 and rel_err(fast.profile, slow.profile) < 1e-12
        assert rel_err(fast.trigger_summary, slow.trigger_summary) < 1e-9


def test_panel_export_paths_agree(ctx):
    """The per-sim panels through the three export paths of lsmc_get_results — page-locked destination (copy engine, all
    panels in flight), pageable destination above the staging threshold (two 32 MB page-locked chunks) and below it —
    are the same bytes."""
    pb = build_config("c3", num_sims=30_000)                       # 366 x 30 000 doubles = 88 MB per panel: three staging chunks
    names = ("inventory", "inject_withdraw", "pv", "spot_valuation", "net_volume")
    pageable = ctx.calculate(pb, Rng.philox(11), Rng.philox(22), panels=names)
    pinned = _abi.Results(pb.n_periods, pb.n_sim_periods, 30_000, panels=names, pinned=True)
    ctx.calculate(pb, Rng.philox(11), Rng.philox(22), results=pinned)
    for name in names:
        assert np.array_equal(getattr(pinned, name), getattr(pageable, name)), name
    assert np.array_equal(pinned.pv_by_sim, pageable.pv_by_sim)
    assert pinned.pv.sum(axis=0) == pytest.approx(pinned.pv_by_sim, rel=1e-12)
    small = ctx.calculate(build_config("c3", num_sims=300), Rng.philox(11), Rng.philox(22), panels=names)   # below the staging threshold
    assert small.pv.sum(axis=0) == pytest.approx(small.pv_by_sim, rel=1e-12)


def test_discounted_deltas_identity(ctx):   # LSMCTest:846-869
    pb = reference_problem(num_sims=256, mean_reversion=16.5)
    a = ctx.calculate(pb, Rng.philox(11), Rng.philox(11))
    b = ctx.calculate(pb.copy(discount_deltas=1), Rng.philox(11), Rng.philox(11))
    assert np.array_equal(b.deltas, a.deltas * pb.settle_df)


def test_progress_is_ascending_ends_at_one_and_counts_like_the_reference(ctx):
    pb = reference_problem(num_sims=128)
    seen = []
    ctx.calculate(pb, Rng.philox(1), Rng.philox(2), on_progress=seen.append)
    assert seen[-1] == 1.0 and 0.0 < seen[0] < 1.0                       # LSMCTest:873-895
    assert all(b > a for a, b in zip(seen, seen[1:]))
    assert len(seen) == 2 * (pb.n_periods - 1) + 1                        # test_multi_factor.py:229-230 (245 for 123 periods)


def test_cancellation_maps_to_operation_cancelled(ctx):   # LSMCTest:899-916
    pb = reference_problem(num_sims=128)
    calls = []

    def cancel_after_ten(v):
        calls.append(v)
        return len(calls) < 10

    with pytest.raises(_native.LsmcError) as e:
        ctx.calculate(pb, Rng.philox(1), Rng.philox(2), on_progress=cancel_after_ten)
    assert e.value.status == _abi.LSMC_ECANCELLED and len(calls) == 10
    again = ctx.calculate(pb, Rng.philox(1), Rng.philox(2))               # the context stays usable
    assert np.isfinite(again.npv)


def test_error_statuses(ctx):
    pb = reference_problem(num_sims=64)
    with pytest.raises(_native.LsmcError) as e:   # start inventory cannot be run down to empty by the end (SH.cs:100-103)
        ctx.calculate(reference_problem(num_sims=64, val_date="2020-03-25", inventory=52_000.0), Rng.philox(1), Rng.philox(2))
    assert e.value.status in (_abi.LSMC_ECONSTRAINT, _abi.LSMC_EINVAL)
    with pytest.raises(_native.LsmcError) as e:
        ctx.calculate(pb.copy(n_basis=0, spot_power=np.zeros(0, dtype=np.int32), factor_power=np.zeros((0, 1), dtype=np.int32)),
                      Rng.philox(1), Rng.philox(2))
    assert e.value.status == _abi.LSMC_EINVAL
    with pytest.raises(_native.LsmcError) as e:
        ctx.forward()                                                     # staged calls out of order
    assert e.value.status == _abi.LSMC_EINVAL
    with pytest.raises(_native.LsmcError) as e:
        ctx.simulate(0, Rng(_abi.RNG_NORMALS))                            # normals mode without normals
    assert e.value.status == _abi.LSMC_EINVAL


def test_shapes_outside_the_device_path_are_refused_as_unsupported_not_computed(ctx):
    """The limits include/lsmc_b200.h documents: each must surface as LSMC_EUNSUPPORTED (the shim's
    NotSupportedException) with a message, never as a wrong number or a CUDA launch error."""
    for kw, needle in [(dict(extra_decisions=7), "extra_decisions"),                              # more than 6 extra decisions per side
                       (dict(num_inventory_grid_points=400), "grid")]:                             # ~200 grid rows at most
        with pytest.raises(_native.LsmcError) as e:
            ctx.calculate(reference_problem(num_sims=64, **kw), Rng.philox(1), Rng.philox(2))
        assert e.value.status == _abi.LSMC_EUNSUPPORTED and needle in str(e.value)
    with pytest.raises(_native.LsmcError) as e:                                                    # continue a generator that never ran
        fresh = _native.Context(0)
        try:
            fresh.calculate(reference_problem(num_sims=64), Rng.mt19937_continue(), Rng.mt19937_continue())
        finally:
            fresh.close()
    assert e.value.status == _abi.LSMC_EINVAL


def test_backward_refuses_factors_of_the_valuation_set(ctx):
    """The factor panel is shared by both simulation sets: after the valuation set has been generated a second
    lsmc_backward would regress regression-set spot on valuation-set factors — it must be refused, not computed."""
    pb = reference_problem(num_sims=256)
    ctx.set_problem(pb)
    ctx.simulate(0, Rng.philox(1))
    ctx.backward()
    ctx.simulate(1, Rng.philox(2))
    with pytest.raises(_native.LsmcError) as e:
        ctx.backward()
    assert e.value.status == _abi.LSMC_EINVAL
    ctx.forward()
    first = ctx.get_results()
    ctx.simulate(0, Rng.philox(1))                                        # regenerating the regression set makes it legal again
    ctx.backward()
    ctx.simulate(1, Rng.philox(2))
    ctx.forward()
    assert ctx.get_results().npv == first.npv


def test_kernel_timing_serialises_the_backward_kernels_without_changing_a_bit(ctx):
    """lsmc_set_kernel_timing puts a CUDA event on either side of every sweep launch, which serialises kernels that otherwise
    overlap by programmatic dependent launch (set-up of a kernel under the tail of its predecessor): a read of a predecessor's
    result placed before the dependency wait would show up as a difference between the two modes.  Few paths, so that the
    grids are smaller than the GPU and several kernels of the chain are resident at once."""
    pb = build_config("c3", num_sims=2048)
    free = ctx.calculate(pb, Rng.philox(5), Rng.philox(6), panels=("inject_withdraw",))
    assert ctx.timings()["backward_sweep_kernel_ms"] == 0.0
    ctx.set_kernel_timing(True)
    try:
        timed = ctx.calculate(pb, Rng.philox(5), Rng.philox(6), panels=("inject_withdraw",))
        tm = ctx.timings()
    finally:
        ctx.set_kernel_timing(False)
    assert tm["backward_sweep_kernel_ms"] > 0.0 and tm["backward_sweep_launches"] == pb.n_periods
    assert timed.npv == free.npv and timed.npv_std_error == free.npv_std_error
    assert np.array_equal(timed.deltas, free.deltas) and np.array_equal(timed.inject_withdraw, free.inject_withdraw)
    assert np.array_equal(timed.trigger_summary, free.trigger_summary, equal_nan=True)


def test_smoke_entry_point():
    import __graft_entry__
    __graft_entry__.smoke()
