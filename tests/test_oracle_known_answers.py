"""Pins the CPU oracle against the reference's own known-answer unit tests (ported; file:line cited per test).

CPU only.  These are what make the oracle a trustworthy checker for the CUDA path (tests/test_gpu_parity.py)."""
import math

import numpy as np
import pytest

from storage_b200._abi import Problem, CONSTRAINT_CONSTANT, CONSTRAINT_PIECEWISE_LINEAR, CONSTRAINT_STEP
from oracle import oracle
from oracle.oracle import OracleError

TOL = 1e-10   # StorageHelperTest.cs:34


# ---- StorageHelper.CalculateBangBangDecisionSet: StorageHelperTest.cs:36-260 ----------------------------------------
BANG_BANG_CASES = [
    # (rmin, rmax, inventory, loss, next_min, next_max, extra, expected builder)   -- StorageHelperTest.cs line
    ((-15.5, 65.685, 1010.0, 10.0, 900.0, 1070.0, 0), lambda lo, hi: [lo, 0.0, hi]),                           # :38
    ((-15.5, 65.685, 1010.0, 10.0, 900.0, 1070.0, 1), lambda lo, hi: [lo, lo / 2.0, 0.0, hi / 2.0, hi]),       # :55
    ((15.5, 65.685, 1010.0, 10.0, 900.0, 1070.0, 0), lambda lo, hi: [lo, hi]),                                 # :74
    ((15.5, 65.685, 1010.0, 10.0, 900.0, 1070.0, 1), lambda lo, hi: [lo, (lo + hi) / 2.0, hi]),                # :91
    ((-65.685, -41.5, 1000.0, 10.0, 900.0, 950.0, 0), lambda lo, hi: [lo, hi]),                                # :109
    ((-65.685, -41.5, 1000.0, 10.0, 900.0, 950.0, 1), lambda lo, hi: [lo, (lo + hi) / 2.0, hi]),               # :126
]


@pytest.mark.parametrize("args,expect", BANG_BANG_CASES)
def test_bang_bang_unconstrained(args, expect):
    rmin, rmax, inv, loss, nmin, nmax, extra = args
    got = oracle.bang_bang_decision_set(rmin, rmax, inv, loss, nmin, nmax, TOL, extra)
    assert got.tolist() == expect(rmin, rmax)


def test_bang_bang_next_step_constrains_both_sides():   # StorageHelperTest.cs:146-184
    inv, loss, nmin, nmax = 1010.0, 10.0, 991.87, 1051.8
    w, j = nmin - inv + loss, nmax - inv + loss   # reference computes (nextMin - (inv - loss))
    got0 = oracle.bang_bang_decision_set(-15.5, 65.685, inv, loss, nmin, nmax, TOL, 0)
    assert got0.tolist() == [nmin - (inv - loss), 0.0, nmax - (inv - loss)]
    np.testing.assert_allclose(got0, [w, 0.0, j], rtol=0, atol=1e-12)
    got1 = oracle.bang_bang_decision_set(-15.5, 65.685, inv, loss, nmin, nmax, TOL, 1)
    np.testing.assert_allclose(got1, [w, w / 2.0, 0.0, j / 2.0, j], rtol=0, atol=1e-12)


def test_bang_bang_injection_lower_than_current():   # StorageHelperTest.cs:186-224
    inv, loss, nmin, nmax = 1010.0, 10.0, 900.0, 995.8
    j = nmax - (inv - loss)
    assert oracle.bang_bang_decision_set(-15.5, 65.685, inv, loss, nmin, nmax, TOL, 0).tolist() == [-15.5, j]
    got = oracle.bang_bang_decision_set(-15.5, 65.685, inv, loss, nmin, nmax, TOL, 1)
    np.testing.assert_allclose(got, [-15.5, (-15.5 + j) / 2.0, j], rtol=0, atol=1e-12)


def test_bang_bang_withdrawal_higher_than_current():   # StorageHelperTest.cs:226-259
    inv, loss, nmin, nmax = 1010.0, 10.0, 1001.8, 1009.51
    w, j = nmin - (inv - loss), nmax - (inv - loss)
    assert oracle.bang_bang_decision_set(-15.5, 65.685, inv, loss, nmin, nmax, TOL, 0).tolist() == [w, j]
    got = oracle.bang_bang_decision_set(-15.5, 65.685, inv, loss, nmin, nmax, TOL, 1)
    np.testing.assert_allclose(got, [w, (w + j) / 2.0, j], rtol=0, atol=1e-12)


def test_bang_bang_argument_checks():   # SH.cs:113-116,130,154
    with pytest.raises(OracleError):
        oracle.bang_bang_decision_set(-1.0, 1.0, 10.0, 0.0, 5.0, 4.0, TOL, 0)      # nextMin > nextMax
    with pytest.raises(OracleError):
        oracle.bang_bang_decision_set(-1.0, 1.0, 10.0, 0.0, 0.0, 5.0, TOL, 0)      # cannot get below nextMax
    with pytest.raises(OracleError):
        oracle.bang_bang_decision_set(-1.0, 1.0, 10.0, 0.0, 20.0, 25.0, TOL, 0)    # cannot reach nextMin
    # within tolerance: clipped to the bound instead of throwing (SH.cs:124-128)
    got = oracle.bang_bang_decision_set(-1.0, 1.0, 10.0, 0.0, 0.0, 9.0 - 1e-12, 1e-10, 0)
    assert got[0] == (9.0 - 1e-12) - 10.0


def test_max_value_and_index():   # StorageHelperTest.cs:262-271
    assert oracle.max_value_and_index([4.5, -1.2, 6.8, 3.2]) == (6.8, 2)
    assert oracle.max_value_and_index([1.0, 7.0, 7.0]) == (7.0, 1)   # first of equal maxima (SH.cs:213: strict >)


# ---- StorageHelper.BisectInventorySpace: StorageHelperTest.cs:429-608 ------------------------------------------------
GRID5 = [0.0, 5.3, 9.5, 15.63, 25.8]
BISECT_CASES = [
    (GRID5, 25.8, (3, 4)),                 # :431 equals highest -> (top-1, top)
    (GRID5, 0.0, (0, 0)),                  # :445
    (GRID5, 20.01, (3, 4)),                # :458
    (GRID5, 1.23, (0, 1)),                 # :472
    (GRID5, 5.3, (1, 1)),                  # :486
    (GRID5, 10.89, (2, 3)),                # :499
    ([1.3, 2.8], 1.4, (0, 1)),             # :512
    ([1.3, 2.8], 1.3, (0, 0)),             # :525
    ([1.3, 2.8], 2.8, (0, 1)),             # :538
    ([1.3], 1.3, (0, 0)),                  # :551
    (GRID5, 0.0 - TOL / 2.0, (0, 0)),      # :584 within tolerance below the minimum
    (GRID5, 25.8 + TOL / 2.0, (4, 4)),     # :597 within tolerance above the maximum
]


@pytest.mark.parametrize("grid,inventory,expect", BISECT_CASES)
def test_bisect_inventory_space(grid, inventory, expect):
    assert oracle.bisect_inventory_space(grid, inventory, TOL) == expect


@pytest.mark.parametrize("inventory", [26.8, -1.2])   # StorageHelperTest.cs:564-582
def test_bisect_outside_grid_throws(inventory):
    with pytest.raises(OracleError):
        oracle.bisect_inventory_space(GRID5, inventory, TOL)


# ---- FixedSpacingStateSpaceGridCalc.GetGridPoints: FixedSpacingStateSpaceGridCalcTest.cs:35-123 ---------------------
def test_grid_points():
    assert oracle.grid_points(120.55, 161.89, 50.0).tolist() == [120.55, 161.89]                 # :36
    assert oracle.grid_points(121.2, 121.2 + 15.0, 15.0).tolist() == [121.2, 121.2 + 15.0]       # :51
    got = oracle.grid_points(121.2, 174.89, 15.0)                                                # :70
    assert got[0] == 121.2 and got[-1] == 174.89 and len(got) == 5
    # xunit's Assert.Equal on double collections passes these literals; accumulated `+=` differs in the last bits
    np.testing.assert_allclose(got[1:4], [136.2, 151.2, 166.2], rtol=0, atol=1e-12)
    assert oracle.grid_points(121.2, 121.2, 15.0).tolist() == [121.2]                            # :93
    with pytest.raises(OracleError):
        oracle.grid_points(121.2, 121.19, 15.0)                                                  # :110


def test_grid_points_accumulate_by_repeated_addition():   # FixedSpacingStateSpaceGridCalc.cs:52-59
    g = oracle.grid_points(0.0, 1.0, 0.1)
    acc, expect = 0.0, [0.0]
    while acc < 1.0:
        acc += 0.1
        expect.append(min(acc, 1.0))
    assert g.tolist() == expect


# ---- constraint problems ---------------------------------------------------------------------------------------------
def constraint_problem(kind, pillars, loss=0.03, min_inv=0.0, max_inv=1000.0, P=3, start_inventory=0.0, must_be_empty=0):
    inv, lo, hi = [np.array(x, dtype=np.float64) for x in zip(*pillars)]
    z, o = np.zeros(P), np.ones(P)
    return Problem(
        n_periods=P, n_factors=1, n_basis=1, n_sims=2, must_be_empty_at_end=must_be_empty, numerical_tolerance=TOL,
        start_inventory=start_inventory, grid_spacing=10.0,
        sim_time=np.arange(1, P + 1) / 365.0, mean_reversion=[1.0], factor_vol=np.full((P, 1), 0.5), factor_corr=[[1.0]],
        forward=np.full(P, 50.0), spot_power=[0], factor_power=[[0]],
        min_inventory=np.full(P, min_inv), max_inventory=np.full(P, max_inv), constraint_id=np.zeros(P, dtype=np.int32),
        inject_cost=z, inject_cost_df=o, withdraw_cost=z, withdraw_cost_df=o, consumed_inject=z, consumed_withdraw=z,
        inventory_loss=np.full(P, loss), inventory_cost=z, inventory_cost_df=o, settle_df=o,
        constraint_kind=[kind], constraint_first=[0], constraint_count=[len(inv)],
        pillar_inventory=inv, pillar_min_rate=lo, pillar_max_rate=hi).finalize()


CONST_RANGE = [(x, -47.12, 56.8) for x in (0.0, 100.0, 300.0, 600.0, 800.0, 1000.0)]
VARYING = [(0.0, -44.85, 56.8), (100.0, -45.01, 54.5), (300.0, -45.78, 52.01), (600.0, -46.17, 51.9), (800.0, -46.99, 50.8),
           (1000.0, -47.12, 50.01)]
STEP_VARYING = VARYING[:-1] + [(1000.0, -46.99, 50.8)]   # StepInjectWithdrawConstraintTest.cs:41-49


@pytest.mark.parametrize("kind", [CONSTRAINT_PIECEWISE_LINEAR, CONSTRAINT_STEP])
def test_space_bounds_constant_rates(kind):
    # PiecewiseLinear…Test.cs:36-64,67-95 and Step…Test.cs:56-83,86-113 (Assert.Equal(..., 12) = 12 decimal places)
    pb = constraint_problem(kind, CONST_RANGE)
    upper, _ = oracle.constraint_bounds(pb, 1, 320.0, 620.0)
    assert abs(upper - (620.0 + 47.12) / (1 - 0.03)) < 0.5e-12 * 1000
    _, lower = oracle.constraint_bounds(pb, 1, 620.0, 870.0)
    assert abs(lower - (620.0 - 56.8) / (1 - 0.03)) < 0.5e-12 * 1000


@pytest.mark.parametrize("kind,pillars,next_hi", [(CONSTRAINT_PIECEWISE_LINEAR, VARYING, 620.0), (CONSTRAINT_STEP, STEP_VARYING, 590.5)])
def test_space_upper_bound_consistent_with_range(kind, pillars, next_hi):
    # PiecewiseLinear…Test.cs:98-125, Step…Test.cs:116-130
    pb = constraint_problem(kind, pillars)
    upper, _ = oracle.constraint_bounds(pb, 1, 320.0, next_hi)
    rmin, _ = oracle.inject_withdraw_range(pb, 1, upper)
    assert abs(upper * (1 - 0.03) + rmin - next_hi) < 1e-11


@pytest.mark.parametrize("kind,pillars", [(CONSTRAINT_PIECEWISE_LINEAR, VARYING), (CONSTRAINT_STEP, STEP_VARYING)])
def test_space_lower_bound_consistent_with_range(kind, pillars):
    # PiecewiseLinear…Test.cs:128-155, Step…Test.cs:133-147
    pb = constraint_problem(kind, pillars)
    _, lower = oracle.constraint_bounds(pb, 1, 552.0, 734.0)
    _, rmax = oracle.inject_withdraw_range(pb, 1, lower)
    assert abs(lower * (1 - 0.03) + rmax - 552.0) < 1e-11


def test_piecewise_linear_step_function_approximation():   # PiecewiseLinear…Test.cs:158-187
    pillars = [(0.0, -5810.0, 7632.0), (64000.0, -5810.0, 7632.0), (64260.0, -15288.0, 7632.0), (182070.0, -15288.0, 3816.0),
               (192780.0, -15288.0, 2279.0), (214200.0, -15288.0, 2269.0)]
    pb = constraint_problem(CONSTRAINT_PIECEWISE_LINEAR, pillars, max_inv=214200.0)
    upper, _ = oracle.constraint_bounds(pb, 1, 320.0, 65260.0)
    rmin, _ = oracle.inject_withdraw_range(pb, 1, upper)
    assert abs(upper * (1 - 0.03) + rmin - 65260.0) < 1e-7   # reference: 12 dp on a value ~6.5e4 with a steep segment


@pytest.mark.parametrize("inventory,rmin,rmax", [(99.0, -44.85, 56.8), (610.85, -46.17, 51.9), (999.99, -46.99, 50.8)])
def test_step_range_lookup(inventory, rmin, rmax):   # Step…Test.cs:149-158
    pb = constraint_problem(CONSTRAINT_STEP, STEP_VARYING)
    assert oracle.inject_withdraw_range(pb, 0, inventory) == (rmin, rmax)


def test_step_range_at_pillars_and_outside():   # Step…Test.cs:161-181
    pb = constraint_problem(CONSTRAINT_STEP, STEP_VARYING)
    for inv, lo, hi in STEP_VARYING:
        assert oracle.inject_withdraw_range(pb, 0, inv) == (lo, hi)
    with pytest.raises(OracleError):
        oracle.inject_withdraw_range(pb, 0, -0.01)
    with pytest.raises(OracleError):
        oracle.inject_withdraw_range(pb, 0, 1000.001)


def test_piecewise_linear_range_at_pillars_and_between():   # PiecewiseLinear…cs:62-72 (LinearSpline)
    pb = constraint_problem(CONSTRAINT_PIECEWISE_LINEAR, VARYING)
    for inv, lo, hi in VARYING:
        got = oracle.inject_withdraw_range(pb, 0, inv)
        assert abs(got[0] - lo) < 1e-12 and abs(got[1] - hi) < 1e-12
    lo, hi = oracle.inject_withdraw_range(pb, 0, 200.0)
    assert abs(lo - (-45.01 + 100.0 * (-45.78 + 45.01) / 200.0)) < 1e-12
    assert abs(hi - (54.5 + 100.0 * (52.01 - 54.5) / 200.0)) < 1e-12


def test_storage_range_outside_inventory_bounds_throws():   # CmdtyStorage.cs:88-94
    pb = constraint_problem(CONSTRAINT_CONSTANT, [(0.0, -6.0, 5.0)], max_inv=23.5)
    assert oracle.inject_withdraw_range(pb, 0, 10.0) == (-6.0, 5.0)
    assert oracle.inject_withdraw_range(pb, 2, 10.0) == (0.0, 0.0)   # period >= EndPeriod (CmdtyStorage.cs:96-97)
    with pytest.raises(OracleError):
        oracle.inject_withdraw_range(pb, 0, 23.6)
    with pytest.raises(OracleError):
        oracle.inject_withdraw_range(pb, 0, -0.1)


# ---- StorageHelper.CalculateInventorySpace: StorageHelperTest.cs:280-427 ---------------------------------------------
def test_inventory_space_current_period_after_storage_start():   # :282-348
    inj, wdr, start_inv, loss, mn, mx = 5.0, 6.0, 8.0, 0.03, 0.0, 23.5
    P = 9   # 2019-08-20 .. 2019-08-28
    pb = constraint_problem(CONSTRAINT_CONSTANT, [(0.0, -wdr, inj)], loss=loss, min_inv=mn, max_inv=mx, P=P, start_inventory=start_inv)
    lo, hi = oracle.inventory_space(pb)
    assert len(lo) == P - 1
    e_lo, e_hi = start_inv * (1 - loss) - wdr, start_inv * (1 - loss) + inj
    # the reference's forward sweep evaluates `inv - loss*inv + rate` (SH.cs:58-75); the test's `inv*(1-loss) + rate` is
    # asserted with exact Assert.Equal and holds for these inputs to the last bit or one ulp -> compare at 1e-14 rel
    assert abs(lo[0] - e_lo) <= 1e-14 * abs(e_lo) and abs(hi[0] - e_hi) <= 1e-14 * abs(e_hi)
    for i in range(1, P - 1):
        e_lo = max(e_lo * (1 - loss) - wdr, mn)
        e_hi = min(e_hi * (1 - loss) + inj, mx)
        assert abs(lo[i] - e_lo) <= 1e-13 * max(1.0, abs(e_lo)) and abs(hi[i] - e_hi) <= 1e-13 * abs(e_hi)


def test_inventory_space_must_be_empty_backward_reduction():   # :350-420
    inj, wdr, start_inv, loss, mn, mx = 5.0, 6.0, 11.0, 0.03, 0.0, 23.5
    P = 10   # 2019-08-19 .. 2019-08-28
    pb = constraint_problem(CONSTRAINT_CONSTANT, [(0.0, -wdr, inj)], loss=loss, min_inv=mn, max_inv=mx, P=P,
                            start_inventory=start_inv, must_be_empty=1)
    pb.max_inventory[-1] = 0.0   # CmdtyStorage.cs:432-438
    lo, hi = oracle.inventory_space(pb)
    assert len(lo) == P - 1
    e_lo, e_hi = start_inv * (1 - loss) - wdr, start_inv * (1 - loss) + inj
    assert abs(lo[0] - e_lo) <= 1e-14 * abs(e_lo) and abs(hi[0] - e_hi) <= 1e-14 * abs(e_hi)
    for i in range(1, 5):   # 21st .. 24th forward sweep
        e_lo = max(e_lo * (1 - loss) - wdr, mn)
        e_hi = min(e_hi * (1 - loss) + inj, mx)
        assert abs(lo[i] - e_lo) <= 1e-13 * max(1.0, abs(e_lo)) and abs(hi[i] - e_hi) <= 1e-13 * abs(e_hi)
    assert (lo[8], hi[8]) == (0.0, 0.0)   # 28th
    e_hi = 0.0
    for i in (7, 6, 5):   # 27th, 26th, 25th: backward reduction
        e_hi = min((e_hi + wdr) / (1 - loss), mx)
        assert lo[i] == 0.0 and abs(hi[i] - e_hi) <= 1e-13 * e_hi


# ---- basis functions: BasisFunctionsBuilderTest.cs:37-71, PowerMonomialBuilderTest.cs:35-57 -------------------------
def test_design_matrix_monomials():
    from storage_b200.utils import parse_basis_functions
    sp, fp = parse_basis_functions("1 + s + x0 + x1**3 + s**2 * x0 * x1**2 + x0*x1", 2)
    pb = constraint_problem(CONSTRAINT_CONSTANT, [(0.0, -1.0, 1.0)])
    pb = pb.copy(n_factors=2, n_basis=len(sp), spot_power=sp, factor_power=fp, mean_reversion=[1.0, 2.0],
                 factor_vol=np.full((3, 2), 0.5), factor_corr=np.eye(2))
    rs = np.random.default_rng(5)
    spot, fac = rs.uniform(10, 90, 17), rs.standard_normal((2, 17))
    X = oracle.design_matrix(pb, spot, fac)
    expect = np.stack([np.ones(17), spot, fac[0], fac[1] ** 3, spot ** 2 * fac[0] * fac[1] ** 2, fac[0] * fac[1]], axis=1)
    np.testing.assert_allclose(X, expect, rtol=4e-16, atol=0)
    assert np.array_equal(X[:, :3], expect[:, :3])   # powers 0 and 1 are exact


def test_regress_matches_numpy_lstsq():   # LSV.cs:187-199: pinv = R^-1 Q^T; beta = pinv y
    rs = np.random.default_rng(9)
    X = np.stack([np.ones(500), rs.standard_normal(500), rs.standard_normal(500) ** 2], axis=1)
    y = X @ np.array([3.0, -2.0, 0.5]) + 0.1 * rs.standard_normal(500)
    beta, fitted = oracle.regress(X, y)
    ref = np.linalg.lstsq(X, y, rcond=None)[0]
    np.testing.assert_allclose(beta, ref, rtol=1e-11)
    np.testing.assert_allclose(fitted, X @ ref, rtol=1e-11)


def test_det_exp_accuracy():
    xs = np.concatenate([np.linspace(-30, 30, 2001), [0.0, 1e-300, -745.0, 709.5, 710.0]])
    for x in xs:
        got, ref = oracle.det_exp(x), math.exp(x) if x < 709.7 else math.inf
        if ref < 1e-300 or math.isinf(ref) or x > 709.0:   # flushes to 0 below exp(-708), saturates to +inf above exp(709)
            assert got == ref or got < 1e-300 or (x > 709.0 and math.isinf(got))
        else:
            assert abs(got / ref - 1.0) < 4e-16


# ---- PolynomialInjectWithdrawConstraint (PolynomialInjectWithdrawConstraint.cs:66-156; C# builder only) ------------------------
def _polynomial_problem(rows, **kw):
    from storage_b200 import CmdtyStorage, RatchetInterp
    from fixtures import reference_problem, STORAGE_START, STORAGE_END
    storage = CmdtyStorage("D", STORAGE_START, STORAGE_END, 1.25, 0.93, ratchets=[(STORAGE_START, rows)],
                           ratchet_interp=RatchetInterp.POLYNOMIAL)
    return reference_problem(storage=storage, num_sims=8, numerical_tolerance=1e-9, **kw), storage


def test_polynomial_constraint_passes_through_its_pillars_and_is_evaluated_by_horner():
    rows = [(0.0, -30.0, 25.0), (500.0, -36.0, 21.0), (1000.0, -38.0, 19.5), (1400.0, -39.0, 19.0)]
    pb, storage = _polynomial_problem(rows, inventory=300.0)
    from storage_b200._abi import CONSTRAINT_POLYNOMIAL
    assert list(pb.constraint_kind) == [CONSTRAINT_POLYNOMIAL] and list(pb.constraint_count) == [4]
    for inv, lo, hi in rows:     # exact-degree fit: the polynomial interpolates the pillars (Fit.Polynomial at order n - 1)
        rmin, rmax = oracle.inject_withdraw_range(pb, 0, inv)
        assert rmin == pytest.approx(lo, abs=1e-6) and rmax == pytest.approx(hi, abs=1e-6)
    coef_lo, coef_hi = pb.pillar_min_rate, pb.pillar_max_rate
    for inv in (123.4, 777.0, 1399.0):   # Polynomial.Evaluate: Horner, each multiply and add rounded
        want_lo, want_hi = coef_lo[-1], coef_hi[-1]
        for k in (2, 1, 0):
            want_lo = want_lo * inv + coef_lo[k]
            want_hi = want_hi * inv + coef_hi[k]
        assert oracle.inject_withdraw_range(pb, 0, inv) == (want_lo, want_hi)
        host = storage.inject_withdraw_range(pb.periods[0], inv)
        assert (host.min_inject_withdraw_rate, host.max_inject_withdraw_rate) == (want_lo, want_hi)


def test_polynomial_table_with_two_equal_rows_collapses_to_constant():   # CmdtyStorageBuilderExtensions.cs:163-171
    pb, _ = _polynomial_problem([(0.0, -20.0, 15.0), (1000.0, -20.0, 15.0)], inventory=100.0)
    from storage_b200._abi import CONSTRAINT_CONSTANT
    assert list(pb.constraint_kind) == [CONSTRAINT_CONSTANT]
    assert oracle.inject_withdraw_range(pb, 3, 400.0) == (-20.0, 15.0)


def test_polynomial_inventory_space_bounds_solve_the_reachability_equation():   # Polynomial…cs:87-156
    rows = [(0.0, -30.0, 25.0), (500.0, -36.0, 21.0), (1000.0, -38.0, 19.5), (1400.0, -39.0, 19.0)]
    pb, _ = _polynomial_problem(rows, inventory=300.0)
    # next period's space [0, 100]: from 1400 one cannot get down to 100 in a day -> the upper bound x solves x + rmin(x) = 100
    upper, lower = oracle.constraint_bounds(pb, 5, 0.0, 100.0)
    rmin, _ = oracle.inject_withdraw_range(pb, 5, upper)
    assert upper + rmin == pytest.approx(100.0, abs=1e-9) and lower == 0.0
    # next period's space [1350, 1400]: the lower bound x solves x + rmax(x) = 1350
    upper, lower = oracle.constraint_bounds(pb, 5, 1350.0, 1400.0)
    _, rmax = oracle.inject_withdraw_range(pb, 5, lower)
    assert lower + rmax == pytest.approx(1350.0, abs=1e-9) and upper == 1400.0
