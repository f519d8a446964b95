/*
 * lsmc_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code).
 *
 * A restatement, in plain double-precision C++, of the Least-Squares Monte Carlo hot path of
 * Cmdty.Storage.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; nothing under storage_b200/ links, imports or calls it.
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference):
 *   LSV.cs  = src/Cmdty.Storage/LsmcValuation/LsmcStorageValuation.cs
 *   LVP.cs  = src/Cmdty.Storage/LsmcValuation/LsmcValuationParameters.cs
 *   SH.cs   = src/Cmdty.Storage/StorageHelper.cs
 *
 * PARITY STATUS
 *  - Engine (backward induction, forward simulation, deltas, trigger prices) and helpers (inventory
 *    space, grids, bang-bang decision sets, bisect, constraints): restated from in-tree C# and PINNED by
 *    the reference's own known-answer tests ported in tests/test_oracle_known_answers.py
 *    (StorageHelperTest.cs, FixedSpacingStateSpaceGridCalcTest.cs, constraint tests, LSMC edge cases
 *    LsmcStorageValuationTest.cs:206-301, Black-76 agreement :309-418, discounted-delta identity :846-869,
 *    trigger-volume identities :984-1016).
 *  - Multi-factor spot simulator: the arithmetic lives in the third-party NuGet package
 *    Cmdty.Core 0.1.0-alpha2020101802 (src/Cmdty.Storage/Cmdty.Storage.csproj:19), whose source is NOT in
 *    /root/reference.  The restatement below follows the model published in-tree
 *    (multi_factor.py:166-187,225-229,258-289; TestHelper.cs:77-86) — "PARITY UNPINNED" at bit level
 *    (draw order / normal transform of MersenneTwisterGenerator are not visible); pinned in distribution
 *    by tests/test_oracle_simulator.py.  Parity runs inject the normals (or whole panels), so the
 *    valuation parity contract does not depend on it.
 *  - Regression: MathNet.Numerics thin QR (LSV.cs:187-190) is third-party; restated as Householder QR,
 *    pseudo-inverse R^-1 Q^T, exactly as the call sequence in the reference.
 *
 * Arithmetic conventions shared with the CUDA kernels so injected-normal runs can be compared bit-wise:
 *  - no fused multiply-add except where written explicitly with fma() (compile with -ffp-contract=off);
 *  - x^q for integer q is a left-to-right product chain (Math.Pow in the reference, BasisFunctions.cs:54,66,111,124;
 *    <= 2 ulp apart);
 *  - exp() in the simulator is det_exp() below (same operations on host and device);
 *  - fitted continuation value X[n,:]·beta (designMatrix.Multiply, LSV.cs:199,400) is an fma chain over k ascending.
 */
#include "../include/lsmc_b200.h"

#include <algorithm>
#include <cmath>
#include <functional>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct OracleError : std::runtime_error {
    int status;
    OracleError(int s, const std::string& m) : std::runtime_error(m), status(s) {}
};

thread_local std::string g_last_error;

// ---------------------------------------------------------------------------------------------
// Deterministic exp (identical operation sequence in storage_b200/csrc/lsmc_device.cuh::det_exp).
// ---------------------------------------------------------------------------------------------
inline double det_exp(double x) {
    if (x > 709.0) return INFINITY;
    if (x < -708.0) return 0.0;
    const double LOG2E = 1.4426950408889634074;
    const double LN2_HI = 6.93147180369123816490e-01;
    const double LN2_LO = 1.90821492927058770002e-10;
    double kf = __builtin_rint(x * LOG2E);
    double r = __builtin_fma(-kf, LN2_HI, x);
    r = __builtin_fma(-kf, LN2_LO, r);
    // Taylor coefficients 1/n!, n = 13 .. 2
    double p = 1.6059043836821613e-10;           // 1/13!
    p = __builtin_fma(p, r, 2.08767569878681e-09);   // 1/12!
    p = __builtin_fma(p, r, 2.505210838544172e-08);  // 1/11!
    p = __builtin_fma(p, r, 2.755731922398589e-07);  // 1/10!
    p = __builtin_fma(p, r, 2.7557319223985893e-06); // 1/9!
    p = __builtin_fma(p, r, 2.48015873015873e-05);   // 1/8!
    p = __builtin_fma(p, r, 0.0001984126984126984);  // 1/7!
    p = __builtin_fma(p, r, 0.001388888888888889);   // 1/6!
    p = __builtin_fma(p, r, 0.008333333333333333);   // 1/5!
    p = __builtin_fma(p, r, 0.041666666666666664);   // 1/4!
    p = __builtin_fma(p, r, 0.16666666666666666);    // 1/3!
    p = __builtin_fma(p, r, 0.5);
    p = __builtin_fma(p, r, 1.0);
    p = __builtin_fma(p, r, 1.0);
    int64_t k = (int64_t)kf;
    uint64_t bits = (uint64_t)(k + 1023) << 52;
    double scale;
    std::memcpy(&scale, &bits, sizeof(double));
    return p * scale;
}

// Integer power as a left-to-right product chain.
inline double chain_mul(double acc, bool& started, double x, int q) {
    for (int i = 0; i < q; i++) {
        if (!started) { acc = x; started = true; }
        else acc = acc * x;
    }
    return acc;
}

// BasisFunctions.Generic (BasisFunctions.cs:92-132): spot^sp * prod_f x_f^q_f.
inline double monomial(const lsmc_problem& pb, int k, double spot, const double* fac) {
    bool started = false;
    double acc = 1.0;
    acc = chain_mul(acc, started, spot, pb.spot_power[k]);
    for (int f = 0; f < pb.n_factors; f++)
        acc = chain_mul(acc, started, fac[f], pb.factor_power[k * pb.n_factors + f]);
    return acc;
}

// ---------------------------------------------------------------------------------------------
// Inject/withdraw constraints.
// ---------------------------------------------------------------------------------------------
struct Constraint {
    int kind;
    int n;
    const double* inv;
    const double* rmin;
    const double* rmax;
};

Constraint get_constraint(const lsmc_problem& pb, int p) {
    int id = pb.constraint_id[p];
    if (id < 0 || id >= pb.n_constraints) throw OracleError(LSMC_EINVAL, "constraint_id out of range");
    Constraint c;
    c.kind = pb.constraint_kind[id];
    c.n = pb.constraint_count[id];
    int first = pb.constraint_first[id];
    c.inv = pb.pillar_inventory + first;
    c.rmin = pb.pillar_min_rate + first;
    c.rmax = pb.pillar_max_rate + first;
    return c;
}

// Array.BinarySearch semantics: index of an exact match or the bitwise complement of the first larger element.
int array_binary_search(const double* a, int n, double v) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        int mid = lo + ((hi - lo) >> 1);
        if (a[mid] == v) return mid;
        if (a[mid] < v) lo = mid + 1;
        else hi = mid - 1;
    }
    return ~lo;
}

// MathNet LinearSpline.InterpolateSorted(...).Interpolate(t) as used by
// PiecewiseLinearInjectWithdrawConstraint.cs:62-72: c0 + (t - x_k) * c1 on the left segment.
double linear_spline(const double* x, const double* y, int n, double t) {
    int index = array_binary_search(x, n, t);
    if (index < 0) index = ~index - 1;
    int k = std::min(std::max(index, 0), n - 2);
    double c1 = (y[k + 1] - y[k]) / (x[k + 1] - x[k]);
    return y[k] + (t - x[k]) * c1;
}

// MathNet Polynomial.Evaluate(z) (coefficients ascending): sum = c[n-1]; for i = n-2..0 { sum *= z; sum += c[i]; }.
double polynomial_evaluate(const double* c, int n, double z) {
    double sum = c[n - 1];
    for (int i = n - 2; i >= 0; --i) { sum *= z; sum += c[i]; }
    return sum;
}

// Polynomial.Differentiate().Evaluate(z): coefficients i * c[i].
double polynomial_derivative(const double* c, int n, double z) {
    if (n < 2) return 0.0;
    double sum = (n - 1) * c[n - 1];
    for (int i = n - 2; i >= 1; --i) { sum *= z; sum += i * c[i]; }
    return sum;
}

// Stand-in for MathNet RobustNewtonRaphson.TryFindRoot(f, df, lower, upper, accuracy, maxIterations, subdivision, out root)
// (PARITY UNPINNED: MathNet is not in the tree).  Specification shared with the product's host plan so that both compute
// the same inventory space: bracket by sign change (cutting [lower, upper] into `subdivision` equal parts from the low end
// when the end points do not bracket), then Newton steps from the mid point, replaced by the bisection point whenever a
// step leaves the bracket; stop when |f| or the step falls below `accuracy`.
bool robust_newton_raphson(const std::function<double(double)>& f, const std::function<double(double)>& df, double lower, double upper,
                           double accuracy, int maxIterations, int subdivision, double& root) {
    double fl = f(lower), fu = f(upper);
    if (fl == 0.0) { root = lower; return true; }
    if (fu == 0.0) { root = upper; return true; }
    if ((fl > 0.0) == (fu > 0.0)) {
        double width = (upper - lower) / subdivision;
        double a = lower, fa = fl;
        bool bracketed = false;
        for (int k = 1; k <= subdivision; k++) {
            double b = (k == subdivision) ? upper : lower + k * width;
            double fb = f(b);
            if (fb == 0.0) { root = b; return true; }
            if ((fa > 0.0) != (fb > 0.0)) { lower = a; upper = b; fl = fa; fu = fb; bracketed = true; break; }
            a = b; fa = fb;
        }
        if (!bracketed) return false;
    }
    double x = 0.5 * (lower + upper);
    for (int it = 0; it < maxIterations; it++) {
        double fx = f(x);
        if (std::fabs(fx) < accuracy) { root = x; return true; }
        if ((fx > 0.0) == (fl > 0.0)) { lower = x; fl = fx; } else { upper = x; fu = fx; }
        double slope = df(x);
        double next = (slope != 0.0) ? x - fx / slope : 0.5 * (lower + upper);
        if (!(next > lower && next < upper)) next = 0.5 * (lower + upper);
        if (std::fabs(next - x) < accuracy) { root = next; return true; }
        x = next;
    }
    return false;
}

// IInjectWithdrawConstraint.GetInjectWithdrawRange: Constant…cs:45-48, PiecewiseLinear…cs:67-72, Step…cs:72-79.
void constraint_range(const Constraint& c, double inventory, double& rmin, double& rmax) {
    switch (c.kind) {
        case LSMC_CONSTRAINT_CONSTANT:
            rmin = c.rmin[0]; rmax = c.rmax[0];
            return;
        case LSMC_CONSTRAINT_PIECEWISE_LINEAR:
            rmax = linear_spline(c.inv, c.rmax, c.n, inventory);
            rmin = linear_spline(c.inv, c.rmin, c.n, inventory);
            return;
        case LSMC_CONSTRAINT_STEP: {
            const double slack = 64.0 * 2.220446049250313e-16 * std::max(1.0, std::fabs(c.inv[c.n - 1]));
            if (inventory < c.inv[0] - slack || inventory > c.inv[c.n - 1] + slack)
                throw OracleError(LSMC_EINVAL, "Value of inventory is outside of the step ratchet interval.");
            int s = array_binary_search(c.inv, c.n, inventory);
            int idx = s >= 0 ? s : ~s - 1;
            idx = std::min(std::max(idx, 0), c.n - 1);
            rmin = c.rmin[idx]; rmax = c.rmax[idx];
            return;
        }
        case LSMC_CONSTRAINT_POLYNOMIAL:   // Polynomial…cs:80-85: both rates are polynomials in the inventory
            rmax = polynomial_evaluate(c.rmax, c.n, inventory);
            rmin = polynomial_evaluate(c.rmin, c.n, inventory);
            return;
        default:
            throw OracleError(LSMC_EUNSUPPORTED, "unknown constraint kind");
    }
}

inline double range_slack(const lsmc_problem& pb, double bound) {
    return std::max(pb.numerical_tolerance, 64.0 * 2.220446049250313e-16 * std::max(1.0, std::fabs(bound)));
}

// CmdtyStorage<T>.GetInjectWithdrawRange (CmdtyStorage.cs:86-100).
void storage_range(const lsmc_problem& pb, int p, double inventory, double& rmin, double& rmax) {
    // DEVIATION (documented in DESIGN.md §6): the reference compares with no tolerance (CmdtyStorage.cs:88-94), so a
    // path whose inventory lands a few ulps outside the bound after `inventory + volume - loss` makes it throw.
    // Both the oracle and the CUDA path accept a breach of at most range_slack(); larger breaches still throw.
    double minInv = pb.min_inventory[p];
    if (inventory < minInv - range_slack(pb, minInv)) throw OracleError(LSMC_EINVAL, "Inventory is below minimum allowed value.");
    double maxInv = pb.max_inventory[p];
    if (inventory > maxInv + range_slack(pb, maxInv)) throw OracleError(LSMC_EINVAL, "Inventory above maximum allowed value.");
    if (p >= pb.n_periods - 1) { rmin = 0.0; rmax = 0.0; return; }
    constraint_range(get_constraint(pb, p), inventory, rmin, rmax);
}

// StorageHelper.InterpolateLinearAndSolve (SH.cs:321-330).
double interpolate_linear_and_solve(double x1, double y1, double x2, double y2, double y) {
    double gradient = (y2 - y1) / (x2 - x1);
    double constant = y1 - gradient * x1;
    return (y - constant) / gradient;
}

// IInjectWithdrawConstraint.InventorySpaceUpperBound:
//   Constant…cs:50-56, PiecewiseLinear…cs:74-117, Step…cs:81-122.
double space_upper_bound(const Constraint& c, double nextLo, double nextHi, double curMin, double curMax, double loss) {
    if (c.kind == LSMC_CONSTRAINT_POLYNOMIAL) {   // Polynomial…cs:87-120
        double maxFromMax = curMax + polynomial_evaluate(c.rmax, c.n, curMax);
        double minFromMax = curMax + polynomial_evaluate(c.rmin, c.n, curMax);
        if (minFromMax <= nextHi && nextLo <= maxFromMax) return curMax;
        auto polyToSolve = [&](double inv) { return inv * (1 - loss) + polynomial_evaluate(c.rmin, c.n, inv) - nextHi; };
        auto polyToSolveDeriv = [&](double inv) { return (1 - loss) + polynomial_derivative(c.rmin, c.n, inv); };
        double thisPeriodMax = 0.0;
        if (!robust_newton_raphson(polyToSolve, polyToSolveDeriv, curMin, curMax, 1e-10, 100, 20, thisPeriodMax))
            throw OracleError(LSMC_ECONSTRAINT, "Cannot solve for the current period inventory space upper bound. Try changing Newton Raphson parameters.");
        if (thisPeriodMax < curMin) throw OracleError(LSMC_ECONSTRAINT, "Inventory constraints cannot be satisfied.");
        return std::min(thisPeriodMax, curMax);
    }
    if (c.kind == LSMC_CONSTRAINT_CONSTANT) {
        double solved = (nextHi - c.rmin[0]) / (1 - loss);
        return std::min(solved, curMax);
    }
    double rmin, rmax;
    constraint_range(c, curMax, rmin, rmax);
    double nextMaxFromMax = curMax * (1 - loss) + rmax;
    double nextMinFromMax = curMax * (1 - loss) + rmin;
    if (nextMinFromMax <= nextHi && nextLo <= nextMaxFromMax) return curMax;
    if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
        double bracketUpperInv = c.inv[c.n - 1];
        double bracketUpperAfter = nextMinFromMax;
        for (int i = c.n - 2; i >= 0; i--) {
            double bracketLowerInv = c.inv[i];
            double bracketLowerAfter = bracketLowerInv * (1 - loss) + c.rmin[i];
            if (bracketLowerAfter <= nextHi && nextHi <= bracketUpperAfter)
                return interpolate_linear_and_solve(bracketLowerInv, bracketLowerAfter, bracketUpperInv, bracketUpperAfter, nextHi);
            bracketUpperAfter = bracketLowerAfter;
            bracketUpperInv = bracketLowerInv;
        }
        throw OracleError(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    }
    // Step
    bool found = false;
    double result = 0.0;
    for (int i = 0; i < c.n - 1; i++) {
        double maxWithdraw = c.rmin[i];
        double lowInv = c.inv[i];
        double lowAfter = lowInv * (1 - loss) + maxWithdraw;
        double upInv = c.inv[i + 1];
        double upAfter = upInv * (1 - loss) + maxWithdraw;
        if (lowAfter <= nextHi && nextHi <= upAfter) {
            result = interpolate_linear_and_solve(lowInv, lowAfter, upInv, upAfter, nextHi);
            found = true;
        }
    }
    if (!found) throw OracleError(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    return result;
}

//   Constant…cs:58-65, PiecewiseLinear…cs:119-160, Step…cs:124-166.
double space_lower_bound(const Constraint& c, double nextLo, double nextHi, double curMin, double curMax, double loss) {
    if (c.kind == LSMC_CONSTRAINT_POLYNOMIAL) {   // Polynomial…cs:122-156
        double maxFromMin = curMin + polynomial_evaluate(c.rmax, c.n, curMin);
        double minFromMin = curMin + polynomial_evaluate(c.rmin, c.n, curMin);
        if (minFromMin <= nextHi && nextLo <= maxFromMin) return curMin;
        auto polyToSolve = [&](double inv) { return inv * (1 - loss) + polynomial_evaluate(c.rmax, c.n, inv) - nextLo; };
        auto polyToSolveDeriv = [&](double inv) { return (1 - loss) + polynomial_derivative(c.rmax, c.n, inv); };
        double thisPeriodMin = 0.0;
        if (!robust_newton_raphson(polyToSolve, polyToSolveDeriv, curMin, curMax, 1e-10, 100, 20, thisPeriodMin))
            throw OracleError(LSMC_ECONSTRAINT, "Cannot solve for the current period inventory space lower bound. Try changing Newton Raphson parameters.");
        if (thisPeriodMin > curMax) throw OracleError(LSMC_ECONSTRAINT, "Inventory constraints cannot be satisfied.");
        return std::max(thisPeriodMin, curMin);
    }
    if (c.kind == LSMC_CONSTRAINT_CONSTANT) {
        double solved = (nextLo - c.rmax[0]) / (1 - loss);
        return std::max(solved, curMin);
    }
    double rmin, rmax;
    constraint_range(c, curMin, rmin, rmax);
    double nextMaxFromMin = curMin * (1 - loss) + rmax;
    double nextMinFromMin = curMin * (1 - loss) + rmin;
    if (nextMinFromMin <= nextHi && nextLo <= nextMaxFromMin) return curMin;
    if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
        double bracketLowerInv = c.inv[0];
        double bracketLowerAfter = nextMaxFromMin;
        for (int i = 1; i < c.n; i++) {
            double bracketUpperInv = c.inv[i];
            double bracketUpperAfter = bracketUpperInv * (1 - loss) + c.rmax[i];
            if (bracketLowerAfter <= nextLo && nextLo <= bracketUpperAfter)
                return interpolate_linear_and_solve(bracketLowerInv, bracketLowerAfter, bracketUpperInv, bracketUpperAfter, nextLo);
            bracketLowerAfter = bracketUpperAfter;
            bracketLowerInv = bracketUpperInv;
        }
        throw OracleError(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    }
    bool found = false;
    double result = 0.0;
    for (int i = c.n - 2; i >= 0; i--) {
        double maxInject = c.rmax[i];
        double lowInv = c.inv[i];
        double lowAfter = lowInv * (1 - loss) + maxInject;
        double upInv = c.inv[i + 1];
        double upAfter = upInv * (1 - loss) + maxInject;
        if (lowAfter <= nextLo && nextLo <= upAfter) {
            result = interpolate_linear_and_solve(lowInv, lowAfter, upInv, upAfter, nextLo);
            found = true;
        }
    }
    if (!found) throw OracleError(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    return result;
}

// StorageHelper.CalculateInventorySpace (SH.cs:39-107).  Output index j = period j+1, j = 0..P-2.
void calculate_inventory_space(const lsmc_problem& pb, std::vector<double>& spaceMin, std::vector<double>& spaceMax) {
    int numPeriods = pb.n_periods - 1;   // storage.EndPeriod.OffsetFrom(startActiveStorage)
    std::vector<double> fwdMax(numPeriods), fwdMin(numPeriods), bwdMax(numPeriods), bwdMin(numPeriods);
    double minFwd = pb.start_inventory, maxFwd = pb.start_inventory;
    for (int i = 0; i < numPeriods; i++) {
        double lossPct = pb.inventory_loss[i];
        double rmin, rmax;
        storage_range(pb, i, minFwd, rmin, rmax);
        double injectWithdrawMin = rmin;
        double lossAtMin = lossPct * minFwd;
        double storageMin = pb.min_inventory[i + 1];
        minFwd = std::max(minFwd - lossAtMin + injectWithdrawMin, storageMin);
        fwdMin[i] = minFwd;

        storage_range(pb, i, maxFwd, rmin, rmax);
        double injectWithdrawMax = rmax;
        double lossAtMax = lossPct * maxFwd;
        double storageMax = pb.max_inventory[i + 1];
        maxFwd = std::min(maxFwd - lossAtMax + injectWithdrawMax, storageMax);
        fwdMax[i] = maxFwd;
    }
    bwdMax[numPeriods - 1] = pb.must_be_empty_at_end ? 0 : pb.max_inventory[pb.n_periods - 1];
    bwdMin[numPeriods - 1] = pb.must_be_empty_at_end ? 0 : pb.min_inventory[pb.n_periods - 1];
    for (int i = numPeriods - 2; i >= 0; i--) {
        int period = i + 1;   // periodBackLoop = EndPeriod offset back
        Constraint c = get_constraint(pb, period);
        double lossPct = pb.inventory_loss[period];
        bwdMax[i] = space_upper_bound(c, bwdMin[i + 1], bwdMax[i + 1], pb.min_inventory[period], pb.max_inventory[period], lossPct);
        bwdMin[i] = space_lower_bound(c, bwdMin[i + 1], bwdMax[i + 1], pb.min_inventory[period], pb.max_inventory[period], lossPct);
    }
    spaceMin.resize(numPeriods); spaceMax.resize(numPeriods);
    for (int i = 0; i < numPeriods; i++) {
        double hi = std::min(fwdMax[i], bwdMax[i]);
        double lo = std::max(fwdMin[i], bwdMin[i]);
        if (lo > hi) throw OracleError(LSMC_ECONSTRAINT, "Inventory constraints cannot be fulfilled.");
        spaceMin[i] = lo; spaceMax[i] = hi;
    }
}

// FixedSpacingStateSpaceGridCalc.GetGridPoints (FixedSpacingStateSpaceGridCalc.cs:45-62).
std::vector<double> grid_points(double lo, double hi, double spacing) {
    if (lo > hi) throw OracleError(LSMC_EINVAL, "stateSpaceLowerBound cannot be above stateSpaceUpperBound");
    std::vector<double> g;
    g.push_back(lo);
    if (lo < hi) {
        double gridPoint = lo;
        do {
            gridPoint += spacing;
            g.push_back(std::min(gridPoint, hi));
        } while (gridPoint < hi);
    }
    return g;
}

// StorageHelper.CalculateBangBangDecisionSet (SH.cs:109-204).
int bang_bang(double rmin, double rmax, double currentInventory, double inventoryLoss, double nextMin, double nextMax,
              double tol, int extra, double* out) {
    if (nextMin > nextMax) throw OracleError(LSMC_EINVAL, "nextStepMinInventory cannot be higher than nextStepMaxInventory.");
    if (extra < 0) throw OracleError(LSMC_EINVAL, "numExtraDecisions must be non-negative.");
    double afterLoss = currentInventory - inventoryLoss;
    double afterMaxWithdrawal = rmin + afterLoss;
    double w;
    if (afterMaxWithdrawal > nextMax) {
        if (afterMaxWithdrawal - nextMax < tol) w = nextMax - afterLoss;
        else throw OracleError(LSMC_EINVAL, "Inventory constraints cannot be fulfilled.");
    } else if (afterMaxWithdrawal > nextMin) {
        w = rmin;
    } else {
        w = nextMin - afterLoss;
    }
    double afterMaxInjection = rmax + afterLoss;
    double j;
    if (afterMaxInjection < nextMin) {
        if (nextMin - afterMaxInjection < tol) j = nextMin - afterLoss;
        else throw OracleError(LSMC_EINVAL, "Inventory constraints cannot be fulfilled.");
    } else if (afterMaxInjection < nextMax) {
        j = rmax;
    } else {
        j = nextMax - afterLoss;
    }
    auto populate = [](double mn, double mx, int n, double* dst) {   // SH.cs:199-204
        double increment = (mx - mn) / (n + 1);
        for (int i = 0; i < n; i++) dst[i] = (i + 1) * increment + mn;
    };
    if (w >= 0.0 || j <= 0.0) {
        if (extra > 0) {
            out[0] = w; out[extra + 1] = j;
            populate(w, j, extra, out + 1);
            return extra + 2;
        }
        out[0] = w; out[1] = j;
        return 2;
    }
    if (extra > 0) {
        int n = extra * 2 + 3;
        out[0] = w; out[n - 1] = j;
        populate(w, 0, extra, out + 1);
        out[extra + 1] = 0.0;
        populate(0, j, extra, out + extra + 2);
        return n;
    }
    out[0] = w; out[1] = 0.0; out[2] = j;
    return 3;
}

// StorageHelper.MaxValueAndIndex (SH.cs:206-221).
inline int max_index(const double* a, int n, double& mx) {
    mx = a[0];
    int idx = 0;
    for (int i = 1; i < n; i++) {
        if (a[i] > mx) { mx = a[i]; idx = i; }
    }
    return idx;
}

// StorageHelper.BisectInventorySpace (SH.cs:280-316).
inline bool equals_within_tol(double a, double b, double tol) { return std::fabs(a - b) <= tol; }

void bisect(const double* grid, int n, double inventory, double tol, int& lower, int& upper) {
    if (n == 1 && equals_within_tol(inventory, grid[0], tol)) { lower = 0; upper = 0; return; }
    int lowerIndex = 0;
    int upperIndex = n - 1;
    int topIndex = upperIndex;
    while (upperIndex > lowerIndex) {
        int mid = (lowerIndex + upperIndex) / 2;
        double invMid = grid[mid];
        if (equals_within_tol(inventory, invMid, tol)) { lower = mid; upper = mid; return; }
        if (invMid > inventory) {
            upperIndex = mid;
        } else {
            int midPlus = mid + 1;
            double invMidPlus = grid[midPlus];
            if (inventory <= invMidPlus) { lower = mid; upper = midPlus; return; }
            if (equals_within_tol(inventory, invMidPlus, tol)) { lower = midPlus; upper = midPlus; return; }
            if (midPlus == topIndex) throw OracleError(LSMC_EINVAL, "Inventory is outside of inventoryGrid bounds.");
            lowerIndex = mid;
        }
    }
    throw OracleError(LSMC_EINVAL, "Inventory is outside of inventoryGrid bounds.");
}

// LSV.cs:672-678 with the builder's percent-of-volume closures (CmdtyStorage.cs:449-497).
inline double consumed_on_decision(const lsmc_problem& pb, int p, double d) {
    return d > 0.0 ? pb.consumed_inject[p] * std::fabs(d) : pb.consumed_withdraw[p] * std::fabs(-d);
}

// LSV.cs:680-689 with per-unit cost closures (CmdtyStorage.cs:327-413): Sum(amount * DF(date)).
inline double inject_withdraw_cost_npv(const lsmc_problem& pb, int p, double d) {
    if (d > 0.0) return (pb.inject_cost[p] * d) * pb.inject_cost_df[p];
    return (pb.withdraw_cost[p] * std::fabs(-d)) * pb.withdraw_cost_df[p];
}

// CmdtyInventoryCost (CmdtyStorage.cs:544-567) summed as at LSV.cs:241-242.
inline double inventory_cost_npv(const lsmc_problem& pb, int p, double inventory) {
    if (pb.inventory_cost[p] == 0.0) return 0.0;   // WithNoInventoryCost: empty cash-flow list sums to 0
    return (inventory * pb.inventory_cost[p]) * pb.inventory_cost_df[p];
}

inline double terminal_npv(const lsmc_problem& pb, double price, double inv) {
    if (pb.must_be_empty_at_end) return 0.0;   // default terminal value (CmdtyStorage.cs:427)
    const double* c = pb.terminal_coef;
    return ((c[0] + c[1] * price) + c[2] * inv) + c[3] * (price * inv);
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011) — same integer stream as the device kernel.
// ---------------------------------------------------------------------------------------------
inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)M0 * c[0];
        uint64_t p1 = (uint64_t)M1 * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += W0; k1 += W1;
    }
}

// Two standard normals for (draw index, step, pair); uniforms in (0,1) from 2x32 bits each.
inline void philox_normal_pair(uint64_t seed, uint64_t draw, uint32_t step, uint32_t pair, double& z0, double& z1) {
    uint32_t c[4] = {(uint32_t)draw, (uint32_t)(draw >> 32), step, pair};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    uint64_t a = ((uint64_t)c[0] << 32) | c[1];
    uint64_t b = ((uint64_t)c[2] << 32) | c[3];
    double u1 = ((double)(a >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    double u2 = ((double)(b >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    double rad = std::sqrt(-2.0 * std::log(u1));
    double ang = 6.283185307179586476925 * u2;
    z0 = rad * std::cos(ang);
    z1 = rad * std::sin(ang);
}

// ---------------------------------------------------------------------------------------------
// Multi-factor simulator (Cmdty.Core MultiFactorSpotPriceSimulator; call site LVP.cs:160-178).
// Model: SURVEY Appendix B.  Per-step coefficient tables.
// ---------------------------------------------------------------------------------------------
struct SimTables {
    int S, F, s0;
    std::vector<double> decay;   // [S][F]
    std::vector<double> chol;    // [S][F][F] lower
    std::vector<double> vol;     // [S][F]
    std::vector<double> drift;   // [S]
    std::vector<double> fwd;     // [S]
};

inline double cont_ext(double x, double dt) {   // multi_factor.py:225-229 with c1 = 0, c2 = -dt
    if (x == 0.0) return dt;
    return -std::expm1(-x * dt) / x;
}

SimTables build_sim_tables(const lsmc_problem& pb) {
    SimTables t;
    t.F = pb.n_factors;
    t.s0 = pb.first_period_is_current ? 1 : 0;
    t.S = pb.n_periods - t.s0;
    int F = t.F;
    t.decay.assign((size_t)t.S * F, 0.0);
    t.chol.assign((size_t)t.S * F * F, 0.0);
    t.vol.assign((size_t)t.S * F, 0.0);
    t.drift.assign(t.S, 0.0);
    t.fwd.assign(t.S, 0.0);
    double tPrev = 0.0;
    for (int j = 0; j < t.S; j++) {
        int p = j + t.s0;
        double tj = pb.sim_time[p];
        double dt = tj - tPrev;
        if (!(dt > 0.0)) throw OracleError(LSMC_EINVAL, "sim_time must be strictly increasing and positive");
        tPrev = tj;
        std::vector<double> cov((size_t)F * F);
        for (int a = 0; a < F; a++) {
            t.decay[(size_t)j * F + a] = std::exp(-pb.mean_reversion[a] * dt);
            t.vol[(size_t)j * F + a] = pb.factor_vol[(size_t)p * F + a];
            for (int b = 0; b < F; b++)
                cov[a * F + b] = pb.factor_corr[a * F + b] * cont_ext(pb.mean_reversion[a] + pb.mean_reversion[b], dt);
        }
        // Cholesky (lower) of the step covariance
        double* L = &t.chol[(size_t)j * F * F];
        for (int a = 0; a < F; a++) {
            for (int b = 0; b <= a; b++) {
                double s = cov[a * F + b];
                for (int k = 0; k < b; k++) s -= L[a * F + k] * L[b * F + k];
                if (a == b) {
                    if (!(s > 0.0)) throw OracleError(LSMC_EINVAL, "factor covariance is not positive definite");
                    L[a * F + a] = std::sqrt(s);
                } else {
                    L[a * F + b] = s / L[b * F + b];
                }
            }
        }
        double var = 0.0;
        for (int a = 0; a < F; a++)
            for (int b = 0; b < F; b++)
                var += pb.factor_vol[(size_t)p * F + a] * pb.factor_vol[(size_t)p * F + b] * pb.factor_corr[a * F + b] *
                       cont_ext(pb.mean_reversion[a] + pb.mean_reversion[b], tj);
        t.drift[j] = 0.5 * var;
        t.fwd[j] = pb.forward[p];
    }
    return t;
}

struct SimSet {
    int S = 0, F = 0;
    int64_t N = 0;
    std::vector<double> spot;   // [S][N]
    std::vector<double> fac;    // [S][F][N]
    const double* spot_row(int j) const { return &spot[(size_t)j * N]; }
    const double* fac_row(int j, int f) const { return &fac[((size_t)j * F + f) * N]; }
};

void simulate(const lsmc_problem& pb, const SimTables& t, const lsmc_rng& rng, SimSet& out) {
    int S = t.S, F = t.F;
    int64_t N = pb.n_sims;
    out.S = S; out.F = F; out.N = N;
    out.spot.assign((size_t)S * N, 0.0);
    out.fac.assign((size_t)S * F * N, 0.0);
    if (rng.mode == LSMC_RNG_PANELS) {
        if (!rng.spot || !rng.factors) throw OracleError(LSMC_EINVAL, "panels mode needs spot and factors");
        std::memcpy(out.spot.data(), rng.spot, sizeof(double) * (size_t)S * N);
        std::memcpy(out.fac.data(), rng.factors, sizeof(double) * (size_t)S * F * N);
        return;
    }
    if (rng.mode == LSMC_RNG_NORMALS && !rng.normals) throw OracleError(LSMC_EINVAL, "normals mode needs normals");
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; n++) {
        double x[LSMC_MAX_FACTORS] = {0};
        double z[LSMC_MAX_FACTORS + 1];
        int64_t gn = pb.sim_offset + n;
        double sign = 1.0;
        uint64_t draw = (uint64_t)gn;
        if (rng.mode == LSMC_RNG_PHILOX && rng.antithetic) { draw = (uint64_t)(gn >> 1); sign = (gn & 1) ? -1.0 : 1.0; }   // partners are adjacent paths
        for (int j = 0; j < S; j++) {
            if (rng.mode == LSMC_RNG_NORMALS) {
                for (int f = 0; f < F; f++) z[f] = rng.normals[((size_t)j * F + f) * N + n];
            } else {
                for (int f = 0; f < F; f += 2) {
                    double a, b;
                    philox_normal_pair(rng.seed, draw, (uint32_t)j, (uint32_t)(f / 2), a, b);
                    z[f] = sign * a; z[f + 1] = sign * b;
                }
            }
            const double* L = &t.chol[(size_t)j * F * F];
            double acc = 0.0;
            for (int f = 0; f < F; f++) {
                double eps = L[f * F + 0] * z[0];
                for (int b = 1; b <= f; b++) eps = eps + L[f * F + b] * z[b];
                x[f] = t.decay[(size_t)j * F + f] * x[f] + eps;
            }
            for (int f = 0; f < F; f++) {
                double term = t.vol[(size_t)j * F + f] * x[f];
                acc = (f == 0) ? term : acc + term;
                out.fac[((size_t)j * F + f) * N + n] = x[f];
            }
            out.spot[(size_t)j * N + n] = t.fwd[j] * det_exp(acc - t.drift[j]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Regression: designMatrix.QR(QRMethod.Thin); R.Inverse(); Q.Transpose(); rInverse * qTranspose
// (LSV.cs:187-190).  Householder QR, explicit thin Q, upper-triangular inverse, pseudo-inverse K x N.
// X is column-major N x K (LSV.cs:766).
// ---------------------------------------------------------------------------------------------
void thin_qr_pseudo_inverse(const std::vector<double>& X, int64_t N, int K, std::vector<double>& pinv /*K x N row-major*/) {
    std::vector<double> A(X);                 // working copy, column-major
    std::vector<double> tau(K, 0.0);
    std::vector<double> R((size_t)K * K, 0.0);
    for (int k = 0; k < K; k++) {
        double* col = &A[(size_t)k * N];
        double normsq = 0.0;
        for (int64_t i = k; i < N; i++) normsq += col[i] * col[i];
        double norm = std::sqrt(normsq);
        double alpha = col[k];
        double beta = alpha >= 0.0 ? -norm : norm;
        if (norm == 0.0) { tau[k] = 0.0; R[(size_t)k * K + k] = 0.0; continue; }
        tau[k] = (beta - alpha) / beta;
        double scale = 1.0 / (alpha - beta);
        for (int64_t i = k + 1; i < N; i++) col[i] *= scale;
        col[k] = 1.0;
        // apply H = I - tau v v^T to the trailing columns
        for (int j = k + 1; j < K; j++) {
            double* cj = &A[(size_t)j * N];
            double dot = 0.0;
            for (int64_t i = k; i < N; i++) dot += col[i] * cj[i];
            dot *= tau[k];
            for (int64_t i = k; i < N; i++) cj[i] -= dot * col[i];
        }
        R[(size_t)k * K + k] = beta;
        for (int j = k + 1; j < K; j++) R[(size_t)k * K + j] = A[(size_t)j * N + k];
    }
    // explicit thin Q = H_0 ... H_{K-1} [I; 0], column-major N x K
    std::vector<double> Q((size_t)N * K, 0.0);
    for (int k = 0; k < K; k++) Q[(size_t)k * N + k] = 1.0;
    for (int k = K - 1; k >= 0; k--) {
        if (tau[k] == 0.0) continue;
        const double* v = &A[(size_t)k * N];
        for (int j = k; j < K; j++) {
            double* qj = &Q[(size_t)j * N];
            double dot = 0.0;
            for (int64_t i = k; i < N; i++) dot += v[i] * qj[i];
            dot *= tau[k];
            for (int64_t i = k; i < N; i++) qj[i] -= dot * v[i];
        }
    }
    // R^-1 (upper triangular)
    std::vector<double> Rinv((size_t)K * K, 0.0);
    for (int j = 0; j < K; j++) {
        Rinv[(size_t)j * K + j] = 1.0 / R[(size_t)j * K + j];
        for (int i = j - 1; i >= 0; i--) {
            double s = 0.0;
            for (int k = i + 1; k <= j; k++) s += R[(size_t)i * K + k] * Rinv[(size_t)k * K + j];
            Rinv[(size_t)i * K + j] = -s / R[(size_t)i * K + i];
        }
    }
    // pinv = Rinv * Q^T
    pinv.assign((size_t)K * N, 0.0);
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; n++) {
        for (int i = 0; i < K; i++) {
            double s = 0.0;
            for (int j = i; j < K; j++) s += Rinv[(size_t)i * K + j] * Q[(size_t)j * N + n];
            pinv[(size_t)i * N + n] = s;
        }
    }
}

// Fitted continuation value for one sim: X[n,:]·beta as an fma chain over k ascending from +0.
inline double fitted_value(const double* Xcolmajor, int64_t N, int K, int64_t n, const double* beta) {
    double acc = 0.0;
    for (int k = 0; k < K; k++) acc = __builtin_fma(Xcolmajor[(size_t)k * N + n], beta[k], acc);
    return acc;
}

// PopulateDesignMatrix (LSV.cs:753-770).
void populate_design_matrix(const lsmc_problem& pb, const SimSet& sims, int j, std::vector<double>& X) {
    int64_t N = sims.N;
    int K = pb.n_basis, F = pb.n_factors;
    X.resize((size_t)N * K);
    const double* spot = sims.spot_row(j);
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; n++) {
        double fac[LSMC_MAX_FACTORS];
        for (int f = 0; f < F; f++) fac[f] = sims.fac_row(j, f)[n];
        for (int k = 0; k < K; k++) X[(size_t)k * N + n] = monomial(pb, k, spot[n], fac);
    }
}

inline double average(const double* v, int64_t n) {   // LSV.cs:691-698
    double sum = 0.0;
    for (int64_t i = 0; i < n; i++) sum += v[i];
    return sum / n;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Optional extra outputs for the parity tests.
// ---------------------------------------------------------------------------------------------
extern "C" {

typedef struct oracle_extras {
    int32_t max_grid;            /* in: row capacity of the per-period buffers below */
    int32_t reserved;
    double* inventory_space_min; /* [P] (index 0 = start inventory) */
    double* inventory_space_max; /* [P] */
    int32_t* grid_count;         /* [P] */
    double* grids;               /* [P][max_grid] */
    double* regression_coeffs;   /* [P-1][max_grid][K]; zero where not regressed */
    double* values;              /* [P][max_grid][N] backward value surfaces, or NULL */
    double* factors_regress;     /* [S][F][N] or NULL */
    double* factors_valuation;   /* [S][F][N] or NULL */
    double* current_continuation;/* [max_grid] means at the current period, or NULL */
    double timing_ms[5];         /* out: regression sim, backward, valuation sim, forward, total */
} oracle_extras;

const char* oracle_last_error(void) { return g_last_error.c_str(); }

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

}  // extern "C"

namespace {

double now_ms() {
#ifdef _OPENMP
    return omp_get_wtime() * 1e3;
#else
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
#endif
}

// LsmcStorageValuation.Calculate<T> (LSV.cs:55-617) from "inputs validated" onwards.
void calculate(const lsmc_problem& pb, const lsmc_rng& regRng, const lsmc_rng& valRng, lsmc_results& res, oracle_extras* ex) {
    const int P = pb.n_periods, K = pb.n_basis, F = pb.n_factors;
    const int64_t N = pb.n_sims;
    if (P < 2) throw OracleError(LSMC_EINVAL, "n_periods must be at least 2");
    if (pb.start_inventory < 0) throw OracleError(LSMC_EINVAL, "Inventory cannot be negative.");   // LSV.cs:61
    if (F < 1 || F > LSMC_MAX_FACTORS || K < 1 || K > LSMC_MAX_BASIS) throw OracleError(LSMC_EINVAL, "bad sizes");
    const int s0 = pb.first_period_is_current ? 1 : 0;
    const int maxDec = 2 * pb.extra_decisions + 3;
    double t_start = now_ms();

    // LSV.cs:88
    std::vector<double> spaceMin, spaceMax;
    calculate_inventory_space(pb, spaceMin, spaceMax);   // index j = period j+1
    auto space_of = [&](int p, double& lo, double& hi) { lo = spaceMin[p - 1]; hi = spaceMax[p - 1]; };

    // LSV.cs:100
    SimTables tables = build_sim_tables(pb);
    SimSet regSims;
    double t0 = now_ms();
    simulate(pb, tables, regRng, regSims);
    double t_regsim = now_ms() - t0;
    if (ex && ex->factors_regress) std::memcpy(ex->factors_regress, regSims.fac.data(), sizeof(double) * regSims.fac.size());
    if (res.spot_regress) std::memcpy(res.spot_regress, regSims.spot.data(), sizeof(double) * regSims.spot.size());

    // grids (LSV.cs:105-111, 209-217)
    std::vector<std::vector<double>> grids(P);
    {
        double lo, hi;
        space_of(P - 1, lo, hi);
        grids[P - 1] = grid_points(lo, hi, pb.grid_spacing);
    }

    // Terminal values (LSV.cs:113-128)
    std::vector<std::vector<double>> actualNext(grids[P - 1].size(), std::vector<double>(N));
    const double* endSpot = regSims.spot_row(P - 1 - s0);
    for (size_t i = 0; i < grids[P - 1].size(); i++)
        for (int64_t n = 0; n < N; n++) actualNext[i][n] = terminal_npv(pb, endSpot[n], grids[P - 1][i]);
    if (ex && ex->values) {
        for (size_t i = 0; i < grids[P - 1].size(); i++)
            std::memcpy(ex->values + ((size_t)(P - 1) * ex->max_grid + i) * N, actualNext[i].data(), sizeof(double) * N);
    }

    std::vector<std::vector<double>> regressCoeffs(P - 1);   // [p] -> [G_next][K]
    std::vector<double> currentPeriodContinuation;
    std::vector<double> X, pinv;

    // Backward induction (LSV.cs:166-340)
    t0 = now_ms();
    for (int p = P - 2; p >= 0; p--) {
        const std::vector<double>& nextGrid = grids[p + 1];
        const int Gn = (int)nextGrid.size();
        std::vector<std::vector<double>> regressNext(Gn);
        const bool isCurrent = (p == 0 && s0 == 1);
        if (isCurrent) {   // LSV.cs:171-182
            currentPeriodContinuation.resize(Gn);
            for (int i = 0; i < Gn; i++) {
                double expected = average(actualNext[i].data(), N);
                regressNext[i].assign(N, expected);
                currentPeriodContinuation[i] = expected;
            }
        } else {   // LSV.cs:185-206
            populate_design_matrix(pb, regSims, p - s0, X);
            thin_qr_pseudo_inverse(X, N, K, pinv);
            regressCoeffs[p].assign((size_t)Gn * K, 0.0);
#pragma omp parallel for schedule(dynamic)
            for (int i = 0; i < Gn; i++) {
                double* beta = &regressCoeffs[p][(size_t)i * K];
                const double* y = actualNext[i].data();
                for (int k = 0; k < K; k++) {
                    const double* prow = &pinv[(size_t)k * N];
                    double s = 0.0;
                    for (int64_t n = 0; n < N; n++) s += prow[n] * y[n];
                    beta[k] = s;
                }
                regressNext[i].resize(N);
                for (int64_t n = 0; n < N; n++) regressNext[i][n] = fitted_value(X.data(), N, K, n, beta);
            }
        }

        // LSV.cs:209-217
        if (p == 0) grids[p] = {pb.start_inventory};
        else { double lo, hi; space_of(p, lo, hi); grids[p] = grid_points(lo, hi, pb.grid_spacing); }
        const std::vector<double>& grid = grids[p];
        double nextMin, nextMax;
        space_of(p + 1, nextMin, nextMax);
        const double df = pb.settle_df[p];
        const double* simPrices = isCurrent ? nullptr : regSims.spot_row(p - s0);
        const double fwdPrice = pb.forward[p];

        std::vector<std::vector<double>> actualThis(grid.size(), std::vector<double>(N));
        std::string loopError; int loopStatus = 0;
#pragma omp parallel for schedule(dynamic)
        for (int ii = 0; ii < (int)grid.size(); ii++) {
            try {
                double inventory = grid[ii];
                double rmin, rmax;
                storage_range(pb, p, inventory, rmin, rmax);                                        // LSV.cs:237
                double inventoryLoss = pb.inventory_loss[p] * inventory;                            // LSV.cs:238
                std::vector<double> decisionSet(maxDec);
                int nd = bang_bang(rmin, rmax, inventory, inventoryLoss, nextMin, nextMax, pb.numerical_tolerance,
                                   pb.extra_decisions, decisionSet.data());                         // LSV.cs:239-240
                double inventoryCostNpv = inventory_cost_npv(pb, p, inventory);                     // LSV.cs:241-242
                std::vector<double> costNpv(nd), consumed(nd);
                std::vector<std::vector<double>> regCont(nd), actCont(nd);
                std::vector<const double*> regPtr(nd, nullptr), actPtr(nd, nullptr);
                for (int d = 0; d < nd; d++) {
                    double vol = decisionSet[d];
                    costNpv[d] = inject_withdraw_cost_npv(pb, p, vol);                              // LSV.cs:254
                    consumed[d] = consumed_on_decision(pb, p, vol);                                 // LSV.cs:257
                    double invAfter = inventory + vol - inventoryLoss;                              // LSV.cs:260
                    for (int g = 0; g < Gn; g++) {                                                  // LSV.cs:261-299
                        double nextInv = nextGrid[g];
                        if (std::fabs(nextInv - invAfter) < 1E-8) {
                            regPtr[d] = regressNext[g].data();
                            actPtr[d] = actualNext[g].data();
                            break;
                        }
                        if (nextInv > invAfter) {
                            if (g == 0) throw OracleError(LSMC_EINVAL, "inventory after decision below next grid");
                            double lowerInv = nextGrid[g - 1];
                            double upperInv = nextInv;
                            double space = upperInv - lowerInv;
                            double wl = (upperInv - invAfter) / space;
                            double wu = 1.0 - wl;
                            regCont[d].resize(N); actCont[d].resize(N);
                            const double* rl = regressNext[g - 1].data(); const double* ru = regressNext[g].data();
                            const double* al = actualNext[g - 1].data(); const double* au = actualNext[g].data();
                            for (int64_t n = 0; n < N; n++) {                                       // WeightedAverage LSV.cs:743-751
                                regCont[d][n] = ru[n] * wu + rl[n] * wl;
                                actCont[d][n] = au[n] * wu + al[n] * wl;
                            }
                            regPtr[d] = regCont[d].data();
                            actPtr[d] = actCont[d].data();
                            break;
                        }
                    }
                    if (!regPtr[d]) throw OracleError(LSMC_EINVAL, "inventory after decision above next grid");
                }
                std::vector<double> npvs(nd);
                double* outv = actualThis[ii].data();
                for (int64_t n = 0; n < N; n++) {                                                   // LSV.cs:304-330
                    double s = simPrices ? simPrices[n] : fwdPrice;
                    for (int d = 0; d < nd; d++) {
                        double vol = decisionSet[d];
                        double injectWithdrawNpv = -vol * s * df;
                        double consumedNpv = -consumed[d] * s * df;
                        double immediate = injectWithdrawNpv - costNpv[d] + consumedNpv;
                        double cont = regPtr[d][n];
                        npvs[d] = immediate + cont - inventoryCostNpv;
                    }
                    double best;
                    int idx = max_index(npvs.data(), nd, best);
                    double adjust = -regPtr[idx][n] + actPtr[idx][n];
                    outv[n] = best + adjust;
                }
            } catch (const OracleError& e) {
#pragma omp critical
                { loopError = e.what(); loopStatus = e.status; }
            }
        }
        if (loopStatus) throw OracleError(loopStatus, loopError);
        if (ex && ex->values) {
            for (size_t i = 0; i < grid.size(); i++)
                std::memcpy(ex->values + ((size_t)p * ex->max_grid + i) * N, actualThis[i].data(), sizeof(double) * N);
        }
        actualNext.swap(actualThis);
    }
    double t_backward = now_ms() - t0;
    double backwardNpv = average(actualNext[0].data(), N);   // LSV.cs:589

    // Valuation sims (LSV.cs:346)
    t0 = now_ms();
    SimSet valSims;
    simulate(pb, tables, valRng, valSims);
    double t_valsim = now_ms() - t0;
    if (ex && ex->factors_valuation) std::memcpy(ex->factors_valuation, valSims.fac.data(), sizeof(double) * valSims.fac.size());
    if (res.spot_valuation) std::memcpy(res.spot_valuation, valSims.spot.data(), sizeof(double) * valSims.spot.size());

    // Forward simulation (LSV.cs:351-562)
    t0 = now_ms();
    std::vector<double> inventoryThis(N, pb.start_inventory), inventoryNext(N);
    std::vector<double> volBySim(N), consumedBySim(N), lossBySim(N), netBySim(N), pvPeriod(N), pvBySim(N, 0.0);
    std::vector<int8_t> decisionIdx(N);
    std::vector<double> deltas(P, 0.0), profile((size_t)P * 6, 0.0);
    const double NaN = std::nan("");
    if (res.trigger_summary) std::fill(res.trigger_summary, res.trigger_summary + (size_t)(P - 1) * 4, NaN);
    if (res.trigger_inject) std::fill(res.trigger_inject, res.trigger_inject + (size_t)(P - 1) * LSMC_NUM_TRIGGER_VOLUMES * 2, NaN);
    if (res.trigger_withdraw) std::fill(res.trigger_withdraw, res.trigger_withdraw + (size_t)(P - 1) * LSMC_NUM_TRIGGER_VOLUMES * 2, NaN);

    for (int p = 0; p < P - 1; p++) {
        const std::vector<double>& nextGrid = grids[p + 1];
        const int Gn = (int)nextGrid.size();
        const bool isCurrent = (p == 0 && s0 == 1);
        std::vector<std::vector<double>> regCont(Gn, std::vector<double>(N));
        if (isCurrent) {   // LSV.cs:382-390
            for (int i = 0; i < Gn; i++) std::fill(regCont[i].begin(), regCont[i].end(), currentPeriodContinuation[i]);
        } else {           // LSV.cs:393-401
            populate_design_matrix(pb, valSims, p - s0, X);
#pragma omp parallel for schedule(dynamic)
            for (int i = 0; i < Gn; i++) {
                const double* beta = &regressCoeffs[p][(size_t)i * K];
                for (int64_t n = 0; n < N; n++) regCont[i][n] = fitted_value(X.data(), N, K, n, beta);
            }
        }
        const double df = pb.settle_df[p];
        const double discountForDeltas = pb.discount_deltas ? df : 1.0;
        const double* simPrices = isCurrent ? nullptr : valSims.spot_row(p - s0);
        const double fwdPrice = pb.forward[p];
        double nextMin, nextMax;
        space_of(p + 1, nextMin, nextMax);

        std::string loopError; int loopStatus = 0;
#pragma omp parallel for schedule(static)
        for (int64_t n = 0; n < N; n++) {   // LSV.cs:426-480
            try {
                double s = simPrices ? simPrices[n] : fwdPrice;
                double inventory = inventoryThis[n];
                double rmin, rmax;
                storage_range(pb, p, inventory, rmin, rmax);
                double inventoryLoss = pb.inventory_loss[p] * inventory;
                double decisionSet[64];
                if (maxDec > 64) throw OracleError(LSMC_EUNSUPPORTED, "too many extra decisions");
                int nd = bang_bang(rmin, rmax, inventory, inventoryLoss, nextMin, nextMax, pb.numerical_tolerance,
                                   pb.extra_decisions, decisionSet);
                double inventoryCostNpv = inventory_cost_npv(pb, p, inventory);
                double npvs[64], used[64], imm[64];
                for (int d = 0; d < nd; d++) {
                    double vol = decisionSet[d];
                    double invAfter = inventory + vol - inventoryLoss;
                    double consumedVol = consumed_on_decision(pb, p, vol);
                    double injectWithdrawNpv = -vol * s * df;
                    double consumedNpv = -consumedVol * s * df;
                    double costNpv = inject_withdraw_cost_npv(pb, p, vol);
                    double immediate = injectWithdrawNpv - costNpv + consumedNpv - inventoryCostNpv;   // LSV.cs:454
                    int lo, hi;
                    bisect(nextGrid.data(), Gn, invAfter, pb.numerical_tolerance, lo, hi);          // LSV.cs:722-741
                    double cont;
                    if (lo == hi) cont = regCont[lo][n];
                    else {
                        double lowerInv = nextGrid[lo], upperInv = nextGrid[hi];
                        double wl = (upperInv - invAfter) / (upperInv - lowerInv);
                        double wu = 1.0 - wl;
                        cont = regCont[lo][n] * wl + regCont[hi][n] * wu;
                    }
                    npvs[d] = immediate + cont;
                    used[d] = consumedVol;
                    imm[d] = immediate;
                }
                double best;
                int idx = max_index(npvs, nd, best);
                double optVol = decisionSet[idx];
                inventoryNext[n] = inventory + optVol - inventoryLoss;
                volBySim[n] = optVol;
                consumedBySim[n] = used[idx];
                lossBySim[n] = inventoryLoss;
                netBySim[n] = -optVol - used[idx];
                pvPeriod[n] = imm[idx];
                pvBySim[n] += imm[idx];
                decisionIdx[n] = (int8_t)idx;
            } catch (const OracleError& e) {
#pragma omp critical
                { loopError = e.what(); loopStatus = e.status; }
            }
        }
        if (loopStatus) throw OracleError(loopStatus, loopError);
        double sumSpotTimesVolume = 0.0;   // LSV.cs:471 (sequential over sims)
        for (int64_t n = 0; n < N; n++) {
            double s = simPrices ? simPrices[n] : fwdPrice;
            sumSpotTimesVolume += -(volBySim[n] + consumedBySim[n]) * s;
        }
        double expectedInventory = average(inventoryThis.data(), N);   // LSV.cs:482-484
        double* prof = &profile[(size_t)p * 6];
        prof[0] = expectedInventory;
        prof[1] = average(volBySim.data(), N);
        prof[2] = average(consumedBySim.data(), N);
        prof[3] = average(lossBySim.data(), N);
        prof[4] = average(netBySim.data(), N);
        prof[5] = average(pvPeriod.data(), N);
        deltas[p] = (sumSpotTimesVolume / fwdPrice / N) * discountForDeltas;   // LSV.cs:485-487

        auto copy_row = [&](double* dst, const std::vector<double>& src) {
            if (dst) std::memcpy(dst + (size_t)p * N, src.data(), sizeof(double) * N);
        };
        copy_row(res.inventory, inventoryThis);
        copy_row(res.inject_withdraw, volBySim);
        copy_row(res.consumed, consumedBySim);
        copy_row(res.loss, lossBySim);
        copy_row(res.net_volume, netBySim);
        copy_row(res.pv, pvPeriod);
        if (res.decision_index) std::memcpy(res.decision_index + (size_t)p * N, decisionIdx.data(), (size_t)N);

        // Trigger prices (LSV.cs:492-561)
        if (res.trigger_summary || res.trigger_inject || res.trigger_withdraw) {
            double expLoss = pb.inventory_loss[p] * expectedInventory;
            double rmin, rmax;
            storage_range(pb, p, expectedInventory, rmin, rmax);
            std::vector<double> ds(maxDec);
            int nd = bang_bang(rmin, rmax, expectedInventory, expLoss, nextMin, nextMax, pb.numerical_tolerance,
                               pb.extra_decisions, ds.data());
            auto avg_cont = [&](double invAfter) {   // AverageContinuationValue LSV.cs:700-720
                int lo, hi;
                bisect(nextGrid.data(), Gn, invAfter, pb.numerical_tolerance, lo, hi);
                if (lo == hi) return average(regCont[lo].data(), N);
                double lowerInv = nextGrid[lo], upperInv = nextGrid[hi];
                double wl = (upperInv - invAfter) / (upperInv - lowerInv);
                double wu = 1.0 - wl;
                double sum = 0.0;
                for (int64_t n = 0; n < N; n++) sum += regCont[lo][n] * wl + regCont[hi][n] * wu;
                return sum / N;
            };
            auto trigger_price = [&](double vol, double altCont, double altVol, double altCost, double altConsumed) {   // LSV.cs:619-638
                double invAfter = expectedInventory + vol - expLoss;
                double contChange = avg_cont(invAfter) - altCont;
                double excessVol = vol - altVol;
                double costChange = inject_withdraw_cost_npv(pb, p, vol) - altCost;
                double consumedChange = consumed_on_decision(pb, p, vol) - altConsumed;
                return (contChange - costChange) / (df * (excessVol + consumedChange));
            };
            double maxInject = *std::max_element(ds.begin(), ds.begin() + nd);
            if (maxInject > 0) {
                double alt = INFINITY;   // smallest non-negative decision (LSV.cs:506-509)
                for (int d = 0; d < nd; d++) if (ds[d] >= 0 && ds[d] < alt) alt = ds[d];
                if (maxInject > alt) {
                    double altCont = avg_cont(expectedInventory + alt - expLoss);   // CalcAlternatives LSV.cs:661-670
                    double altCost = inject_withdraw_cost_npv(pb, p, alt);
                    double altConsumed = consumed_on_decision(pb, p, alt);
                    const int nv = LSMC_NUM_TRIGGER_VOLUMES;
                    double vols[LSMC_NUM_TRIGGER_VOLUMES];   // LSV.cs:640-649
                    double inc = (maxInject - alt) / nv;
                    vols[nv - 1] = maxInject;
                    for (int i = 1; i < nv; i++) vols[i - 1] = alt + i * inc;
                    double lastPrice = NaN;
                    for (int i = 0; i < nv; i++) {
                        double price = trigger_price(vols[i], altCont, alt, altCost, altConsumed);
                        if (res.trigger_inject) {
                            res.trigger_inject[((size_t)p * nv + i) * 2] = vols[i];
                            res.trigger_inject[((size_t)p * nv + i) * 2 + 1] = price;
                        }
                        lastPrice = price;
                    }
                    if (res.trigger_summary) { res.trigger_summary[(size_t)p * 4 + 0] = maxInject; res.trigger_summary[(size_t)p * 4 + 1] = lastPrice; }
                }
            }
            double maxWithdraw = *std::min_element(ds.begin(), ds.begin() + nd);
            if (maxWithdraw < 0) {
                double alt = -INFINITY;   // largest non-positive decision (LSV.cs:534-537)
                for (int d = 0; d < nd; d++) if (ds[d] <= 0 && ds[d] > alt) alt = ds[d];
                if (maxWithdraw < alt) {
                    double altCont = avg_cont(expectedInventory + alt - expLoss);
                    double altCost = inject_withdraw_cost_npv(pb, p, alt);
                    double altConsumed = consumed_on_decision(pb, p, alt);
                    const int nv = LSMC_NUM_TRIGGER_VOLUMES;
                    double vols[LSMC_NUM_TRIGGER_VOLUMES];   // LSV.cs:651-659
                    double inc = (alt - maxWithdraw) / nv;
                    for (int i = 0; i < nv; i++) vols[i] = maxWithdraw + i * inc;
                    double firstPrice = NaN;
                    for (int i = nv - 1, outIdx = 0; i >= 0; i--, outIdx++) {   // iterated in reverse (LSV.cs:545)
                        double price = trigger_price(vols[i], altCont, alt, altCost, altConsumed);
                        if (res.trigger_withdraw) {
                            res.trigger_withdraw[((size_t)p * nv + outIdx) * 2] = vols[i];
                            res.trigger_withdraw[((size_t)p * nv + outIdx) * 2 + 1] = price;
                        }
                        if (outIdx == 0) firstPrice = price;
                    }
                    if (res.trigger_summary) { res.trigger_summary[(size_t)p * 4 + 2] = maxWithdraw; res.trigger_summary[(size_t)p * 4 + 3] = firstPrice; }
                }
            }
        }
        inventoryThis.swap(inventoryNext);
    }

    // Final period (LSV.cs:563-579): terminal value on the REGRESSION sims' end prices.
    double endPeriodPv = 0.0;
    if (res.inventory) std::memcpy(res.inventory + (size_t)(P - 1) * N, inventoryThis.data(), sizeof(double) * N);
    auto zero_last = [&](double* dst) { if (dst) std::fill(dst + (size_t)(P - 1) * N, dst + (size_t)P * N, 0.0); };
    zero_last(res.inject_withdraw); zero_last(res.consumed); zero_last(res.loss); zero_last(res.net_volume); zero_last(res.pv);
    if (!pb.must_be_empty_at_end) {
        std::vector<double> endPv(N);
        for (int64_t n = 0; n < N; n++) {
            double tpv = terminal_npv(pb, endSpot[n], inventoryThis[n]);
            endPv[n] = tpv;
            pvBySim[n] += tpv;
        }
        if (res.pv) std::memcpy(res.pv + (size_t)(P - 1) * N, endPv.data(), sizeof(double) * N);
        endPeriodPv = average(endPv.data(), N);
    }
    double t_forward = now_ms() - t0;

    res.npv = average(pvBySim.data(), N);   // LSV.cs:584
    res.backward_npv = backwardNpv;
    {   // standard error from antithetic pair means when the draws are paired, else plain
        bool paired = valRng.mode == LSMC_RNG_PHILOX && valRng.antithetic && (N % 2 == 0) && (pb.sim_offset % 2 == 0);
        int64_t m = paired ? N / 2 : N;
        double mean = 0.0, m2 = 0.0;
        for (int64_t i = 0; i < m; i++) {
            double v = paired ? 0.5 * (pvBySim[2 * i] + pvBySim[2 * i + 1]) : pvBySim[i];
            double dlt = v - mean;
            mean += dlt / (double)(i + 1);
            m2 += dlt * (v - mean);
        }
        res.npv_std_error = m > 1 ? std::sqrt(m2 / (double)(m - 1) / (double)m) : 0.0;
    }
    double expectedFinalInventory = average(inventoryThis.data(), N);   // LSV.cs:593-595
    profile[(size_t)(P - 1) * 6 + 0] = expectedFinalInventory;
    profile[(size_t)(P - 1) * 6 + 5] = endPeriodPv;
    if (res.deltas) std::memcpy(res.deltas, deltas.data(), sizeof(double) * P);
    if (res.profile) std::memcpy(res.profile, profile.data(), sizeof(double) * P * 6);
    if (res.pv_by_sim) std::memcpy(res.pv_by_sim, pvBySim.data(), sizeof(double) * N);

    if (ex) {
        if (ex->inventory_space_min && ex->inventory_space_max) {
            ex->inventory_space_min[0] = pb.start_inventory; ex->inventory_space_max[0] = pb.start_inventory;
            for (int p = 1; p < P; p++) { ex->inventory_space_min[p] = spaceMin[p - 1]; ex->inventory_space_max[p] = spaceMax[p - 1]; }
        }
        for (int p = 0; p < P; p++) {
            if ((int)grids[p].size() > ex->max_grid) throw OracleError(LSMC_EINVAL, "oracle_extras.max_grid too small");
            if (ex->grid_count) ex->grid_count[p] = (int)grids[p].size();
            if (ex->grids) std::memcpy(ex->grids + (size_t)p * ex->max_grid, grids[p].data(), sizeof(double) * grids[p].size());
        }
        if (ex->regression_coeffs) {
            std::fill(ex->regression_coeffs, ex->regression_coeffs + (size_t)(P - 1) * ex->max_grid * K, 0.0);
            for (int p = 0; p < P - 1; p++)
                if (!regressCoeffs[p].empty())
                    std::memcpy(ex->regression_coeffs + (size_t)p * ex->max_grid * K, regressCoeffs[p].data(), sizeof(double) * regressCoeffs[p].size());
        }
        if (ex->current_continuation && !currentPeriodContinuation.empty())
            std::memcpy(ex->current_continuation, currentPeriodContinuation.data(), sizeof(double) * currentPeriodContinuation.size());
        ex->timing_ms[0] = t_regsim; ex->timing_ms[1] = t_backward; ex->timing_ms[2] = t_valsim; ex->timing_ms[3] = t_forward;
        ex->timing_ms[4] = now_ms() - t_start;
    }
}

template <typename Fn>
int guarded(Fn&& fn) {
    try {
        fn();
        return LSMC_OK;
    } catch (const OracleError& e) {
        g_last_error = e.what();
        return e.status;
    } catch (const std::exception& e) {
        g_last_error = e.what();
        return LSMC_EINVAL;
    }
}

}  // namespace

extern "C" {

int oracle_lsmc_calculate(const lsmc_problem* pb, const lsmc_rng* reg, const lsmc_rng* val, lsmc_results* res, oracle_extras* ex) {
    return guarded([&] {
        if (!pb || !reg || !val || !res) throw OracleError(LSMC_EINVAL, "null argument");
        if (pb->struct_size != (int32_t)sizeof(lsmc_problem)) throw OracleError(LSMC_EINVAL, "lsmc_problem.struct_size mismatch");
        calculate(*pb, *reg, *val, *res, ex);
    });
}

/* --- helper entry points for the known-answer tests ported from the reference's unit tests --- */

int oracle_bang_bang_decision_set(double rmin, double rmax, double inventory, double loss, double nextMin, double nextMax,
                                  double tol, int extra, double* out, int* count) {
    return guarded([&] { *count = bang_bang(rmin, rmax, inventory, loss, nextMin, nextMax, tol, extra, out); });
}

int oracle_bisect_inventory_space(const double* grid, int n, double inventory, double tol, int* lower, int* upper) {
    return guarded([&] { int lo, hi; bisect(grid, n, inventory, tol, lo, hi); *lower = lo; *upper = hi; });
}

int oracle_max_value_and_index(const double* a, int n, double* mx, int* idx) {
    return guarded([&] { double m; *idx = max_index(a, n, m); *mx = m; });
}

int oracle_grid_points(double lo, double hi, double spacing, double* out, int cap, int* count) {
    return guarded([&] {
        auto g = grid_points(lo, hi, spacing);
        if ((int)g.size() > cap) throw OracleError(LSMC_EINVAL, "grid buffer too small");
        std::copy(g.begin(), g.end(), out);
        *count = (int)g.size();
    });
}

int oracle_inventory_space(const lsmc_problem* pb, double* mins /*[P-1]*/, double* maxs /*[P-1]*/) {
    return guarded([&] {
        std::vector<double> lo, hi;
        calculate_inventory_space(*pb, lo, hi);
        std::copy(lo.begin(), lo.end(), mins);
        std::copy(hi.begin(), hi.end(), maxs);
    });
}

int oracle_inject_withdraw_range(const lsmc_problem* pb, int period, double inventory, double* rmin, double* rmax) {
    return guarded([&] { storage_range(*pb, period, inventory, *rmin, *rmax); });
}

int oracle_constraint_bounds(const lsmc_problem* pb, int period, double nextLo, double nextHi, double* upper, double* lower) {
    return guarded([&] {
        Constraint c = get_constraint(*pb, period);
        double loss = pb->inventory_loss[period];
        *upper = space_upper_bound(c, nextLo, nextHi, pb->min_inventory[period], pb->max_inventory[period], loss);
        *lower = space_lower_bound(c, nextLo, nextHi, pb->min_inventory[period], pb->max_inventory[period], loss);
    });
}

int oracle_design_matrix(const lsmc_problem* pb, const double* spot, const double* factors /*[F][N]*/, int64_t N, double* X /*col-major N x K*/) {
    return guarded([&] {
        for (int64_t n = 0; n < N; n++) {
            double fac[LSMC_MAX_FACTORS];
            for (int f = 0; f < pb->n_factors; f++) fac[f] = factors[(size_t)f * N + n];
            for (int k = 0; k < pb->n_basis; k++) X[(size_t)k * N + n] = monomial(*pb, k, spot[n], fac);
        }
    });
}

int oracle_simulate(const lsmc_problem* pb, const lsmc_rng* rng, double* spot /*[S][N]*/, double* factors /*[S][F][N]*/) {
    return guarded([&] {
        SimTables t = build_sim_tables(*pb);
        SimSet s;
        simulate(*pb, t, *rng, s);
        if (spot) std::memcpy(spot, s.spot.data(), sizeof(double) * s.spot.size());
        if (factors) std::memcpy(factors, s.fac.data(), sizeof(double) * s.fac.size());
    });
}

int oracle_regress(const double* X /*col-major N x K*/, int64_t N, int K, const double* y, double* beta, double* fitted) {
    return guarded([&] {
        std::vector<double> Xv(X, X + (size_t)N * K), pinv;
        thin_qr_pseudo_inverse(Xv, N, K, pinv);
        for (int k = 0; k < K; k++) {
            double s = 0.0;
            for (int64_t n = 0; n < N; n++) s += pinv[(size_t)k * N + n] * y[n];
            beta[k] = s;
        }
        if (fitted) for (int64_t n = 0; n < N; n++) fitted[n] = fitted_value(X, N, K, n, beta);
    });
}

double oracle_det_exp(double x) { return det_exp(x); }

/* Philox4x32-10 block function (pinned by the Random123 known-answer vectors in tests/test_oracle_simulator.py). */
void oracle_philox4x32_10(uint32_t ctr[4], uint32_t k0, uint32_t k1) { philox4x32_10(ctr, k0, k1); }

/* The two standard normals the simulators draw for (seed, draw index, step, factor pair). */
void oracle_philox_normal_pair(uint64_t seed, uint64_t draw, uint32_t step, uint32_t pair, double* z0, double* z1) {
    philox_normal_pair(seed, draw, step, pair, *z0, *z1);
}

}  // extern "C"
