// host_plan.cpp — see host_plan.h.  Compile with -ffp-contract=off: every a*b+c below is two rounded
// operations, as in the reference's managed code.
#include "host_plan.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>

namespace lsmc {

// ------------------------------------------------------------------------------------------------
// Problem copy
// ------------------------------------------------------------------------------------------------
template <typename T>
static void take(std::vector<T>& dst, const T*& field, size_t n, const char* name) {
    if (n > 0 && field == nullptr) throw Error(LSMC_EINVAL, std::string("lsmc_problem.") + name + " is null");
    dst.assign(field, field + n);
    field = dst.data();
}

void ProblemCopy::assign(const lsmc_problem& src) {
    if (src.struct_size != (int32_t)sizeof(lsmc_problem))
        throw Error(LSMC_EINVAL, "lsmc_problem.struct_size does not match this library's header");
    pb = src;
    const size_t P = (size_t)src.n_periods, F = (size_t)src.n_factors, K = (size_t)src.n_basis;
    if (src.n_periods < 2) throw Error(LSMC_EINVAL, "n_periods must be at least 2");
    if (src.n_factors < 1 || src.n_factors > LSMC_MAX_FACTORS) throw Error(LSMC_EINVAL, "n_factors out of range");
    if (src.n_basis < 1 || src.n_basis > LSMC_MAX_BASIS) throw Error(LSMC_EINVAL, "n_basis out of range");
    if (src.n_sims < 1) throw Error(LSMC_EINVAL, "n_sims must be positive");
    if (src.extra_decisions < 0) throw Error(LSMC_EINVAL, "ExtraDecisions must be non-negative.");   // LVP.cs:134
    if (src.start_inventory < 0) throw Error(LSMC_EINVAL, "Inventory cannot be negative.");          // LSV.cs:61
    if (!(src.grid_spacing > 0.0)) throw Error(LSMC_EINVAL, "grid spacing must be positive");         // FixedSpacing…cs:40
    if (src.n_constraints < 1 || src.n_pillars < 1) throw Error(LSMC_EINVAL, "no inject/withdraw constraints");
#define TAKE(field, n) take(field, pb.field, (n), #field)
    TAKE(sim_time, P); TAKE(mean_reversion, F); TAKE(factor_vol, P * F); TAKE(factor_corr, F * F); TAKE(forward, P);
    TAKE(spot_power, K); TAKE(factor_power, K * F);
    TAKE(min_inventory, P); TAKE(max_inventory, P); TAKE(constraint_id, P);
    TAKE(inject_cost, P); TAKE(inject_cost_df, P); TAKE(withdraw_cost, P); TAKE(withdraw_cost_df, P);
    TAKE(consumed_inject, P); TAKE(consumed_withdraw, P); TAKE(inventory_loss, P);
    TAKE(inventory_cost, P); TAKE(inventory_cost_df, P); TAKE(settle_df, P);
    const size_t C = (size_t)src.n_constraints, NP = (size_t)src.n_pillars;
    TAKE(constraint_kind, C); TAKE(constraint_first, C); TAKE(constraint_count, C);
    TAKE(pillar_inventory, NP); TAKE(pillar_min_rate, NP); TAKE(pillar_max_rate, NP);
#undef TAKE
    for (size_t k = 0; k < K; k++) {
        if (pb.spot_power[k] < 0) throw Error(LSMC_EINVAL, "Spot power must be non-negative.");   // BasisFunctions.cs:94
        for (size_t f = 0; f < F; f++)
            if (pb.factor_power[k * F + f] < 0) throw Error(LSMC_EINVAL, "Markov factor powers must be non-negative.");
    }
    for (size_t p = 0; p < P; p++)
        if (pb.constraint_id[p] < 0 || pb.constraint_id[p] >= src.n_constraints) throw Error(LSMC_EINVAL, "constraint_id out of range");
    for (size_t c = 0; c < C; c++) {
        int kind = pb.constraint_kind[c], first = pb.constraint_first[c], count = pb.constraint_count[c];
        if (kind < LSMC_CONSTRAINT_CONSTANT || kind > LSMC_CONSTRAINT_POLYNOMIAL) throw Error(LSMC_EUNSUPPORTED, "unsupported inject/withdraw constraint kind");
        if (kind == LSMC_CONSTRAINT_POLYNOMIAL && count > LSMC_MAX_POLYNOMIAL_PILLARS) throw Error(LSMC_EUNSUPPORTED, "polynomial inject/withdraw constraint with more than 16 pillars");
        if (first < 0 || count < 1 || first + count > src.n_pillars) throw Error(LSMC_EINVAL, "constraint pillar range out of bounds");
        if (kind != LSMC_CONSTRAINT_CONSTANT && count < 2)
            throw Error(LSMC_EINVAL, "Inject/withdraw ranges collection must contain at least two elements.");
    }
}

// ------------------------------------------------------------------------------------------------
// Constraints: rate lookup and inventory-space bounds
// ------------------------------------------------------------------------------------------------
namespace {

struct Pillars {
    int kind, n;
    const double *x, *lo, *hi;
};

Pillars pillars_of(const lsmc_problem& pb, int period) {
    int id = pb.constraint_id[period];
    int first = pb.constraint_first[id];
    return Pillars{pb.constraint_kind[id], pb.constraint_count[id], pb.pillar_inventory + first,
                   pb.pillar_min_rate + first, pb.pillar_max_rate + first};
}

// index of the last pillar <= v, or -1 (Array.BinarySearch + complement, Step…cs:76-77 / MathNet LeftSegmentIndex)
int last_pillar_at_or_below(const double* x, int n, double v) {
    int idx = (int)(std::upper_bound(x, x + n, v) - x) - 1;
    return idx;
}

double spline_eval(const double* x, const double* y, int n, double v) {
    int k = std::min(std::max(last_pillar_at_or_below(x, n, v), 0), n - 2);
    double slope = (y[k + 1] - y[k]) / (x[k + 1] - x[k]);
    return y[k] + (v - x[k]) * slope;
}

// Polynomial.Evaluate (MathNet): Horner's rule from the highest coefficient, every multiply and add rounded.
double horner(const double* coef, int n, double z) {
    double sum = coef[n - 1];
    for (int i = n - 2; i >= 0; i--) { sum *= z; sum += coef[i]; }
    return sum;
}

double horner_derivative(const double* coef, int n, double z) {   // Polynomial.Differentiate().Evaluate
    if (n < 2) return 0.0;
    double sum = (n - 1) * coef[n - 1];
    for (int i = n - 2; i >= 1; i--) { sum *= z; sum += i * coef[i]; }
    return sum;
}

// Root of f on [lo, hi] by Newton's method kept inside a sign-change bracket (bisection when a step leaves it); when the
// end points do not bracket, the interval is cut into `subdivision` parts and the lowest part with a sign change is used.
// Stands in for MathNet's RobustNewtonRaphson.TryFindRoot (PolynomialInjectWithdrawConstraint.cs:109,144), which is not in
// the tree: same contract (accuracy 1e-10, 100 iterations, 20 subdivisions by default), not the same iterates.
template <typename Fn, typename Dfn>
bool find_root(const Fn& f, const Dfn& df, double lo, double hi, double accuracy, int max_iterations, int subdivision, double& root) {
    double flo = f(lo), fhi = f(hi);
    if (flo == 0.0) { root = lo; return true; }
    if (fhi == 0.0) { root = hi; return true; }
    if ((flo > 0.0) == (fhi > 0.0)) {
        const double step = (hi - lo) / subdivision;
        bool found = false;
        double a = lo, fa = flo;
        for (int k = 1; k <= subdivision && !found; k++) {
            const double b = k == subdivision ? hi : lo + k * step, fb = f(b);
            if (fb == 0.0) { root = b; return true; }
            if ((fa > 0.0) != (fb > 0.0)) { lo = a; hi = b; flo = fa; fhi = fb; found = true; }
            a = b; fa = fb;
        }
        if (!found) return false;
    }
    double x = 0.5 * (lo + hi);
    for (int it = 0; it < max_iterations; it++) {
        const double fx = f(x);
        if (std::fabs(fx) < accuracy) { root = x; return true; }
        if ((fx > 0.0) == (flo > 0.0)) { lo = x; flo = fx; } else { hi = x; fhi = fx; }
        const double d = df(x);
        double xn = d != 0.0 ? x - fx / d : 0.5 * (lo + hi);
        if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
        if (std::fabs(xn - x) < accuracy) { root = xn; return true; }
        x = xn;
    }
    return false;
}

void rate_range(const Pillars& c, double v, double& rmin, double& rmax) {
    if (c.kind == LSMC_CONSTRAINT_CONSTANT) { rmin = c.lo[0]; rmax = c.hi[0]; return; }
    if (c.kind == LSMC_CONSTRAINT_POLYNOMIAL) { rmax = horner(c.hi, c.n, v); rmin = horner(c.lo, c.n, v); return; }
    if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
        rmax = spline_eval(c.x, c.hi, c.n, v);
        rmin = spline_eval(c.x, c.lo, c.n, v);
        return;
    }
    const double slack = range_slack(0.0, c.x[c.n - 1]);
    if (v < c.x[0] - slack || v > c.x[c.n - 1] + slack) throw Error(LSMC_EINVAL, "Value of inventory is outside of the ratchet interval.");
    int idx = std::min(std::max(last_pillar_at_or_below(c.x, c.n, v), 0), c.n - 1);
    rmin = c.lo[idx]; rmax = c.hi[idx];
}

double solve_on_line(double x1, double y1, double x2, double y2, double y) {   // SH.cs:321-330
    double gradient = (y2 - y1) / (x2 - x1);
    double constant = y1 - gradient * x1;
    return (y - constant) / gradient;
}

// Largest inventory of this period from which [nextLo, nextHi] is reachable.
// PolynomialInjectWithdrawConstraint.InventorySpaceUpperBound / LowerBound (:87-156).  (The reachability test adds the rate to
// the bound WITHOUT the loss, the root equation has it: as the reference writes them.)
double polynomial_upper_bound(const Pillars& c, double nextLo, double nextHi, double curMin, double curMax, double loss) {
    const double reachHi = curMax + horner(c.hi, c.n, curMax), reachLo = curMax + horner(c.lo, c.n, curMax);
    if (reachLo <= nextHi && nextLo <= reachHi) return curMax;
    double root = 0.0;
    auto f = [&](double inv) { return inv * (1 - loss) + horner(c.lo, c.n, inv) - nextHi; };
    auto df = [&](double inv) { return (1 - loss) + horner_derivative(c.lo, c.n, inv); };
    if (!find_root(f, df, curMin, curMax, 1e-10, 100, 20, root))
        throw Error(LSMC_ECONSTRAINT, "Cannot solve for the current period inventory space upper bound. Try changing Newton Raphson parameters.");
    if (root < curMin) throw Error(LSMC_ECONSTRAINT, "Inventory constraints cannot be satisfied.");
    return std::min(root, curMax);
}

double polynomial_lower_bound(const Pillars& c, double nextLo, double nextHi, double curMin, double curMax, double loss) {
    const double reachHi = curMin + horner(c.hi, c.n, curMin), reachLo = curMin + horner(c.lo, c.n, curMin);
    if (reachLo <= nextHi && nextLo <= reachHi) return curMin;
    double root = 0.0;
    auto f = [&](double inv) { return inv * (1 - loss) + horner(c.hi, c.n, inv) - nextLo; };
    auto df = [&](double inv) { return (1 - loss) + horner_derivative(c.hi, c.n, inv); };
    if (!find_root(f, df, curMin, curMax, 1e-10, 100, 20, root))
        throw Error(LSMC_ECONSTRAINT, "Cannot solve for the current period inventory space lower bound. Try changing Newton Raphson parameters.");
    if (root > curMax) throw Error(LSMC_ECONSTRAINT, "Inventory constraints cannot be satisfied.");
    return std::max(root, curMin);
}

double upper_bound_of_space(const Pillars& c, double nextLo, double nextHi, double curMax, double loss) {
    if (c.kind == LSMC_CONSTRAINT_CONSTANT) return std::min((nextHi - c.lo[0]) / (1 - loss), curMax);
    double rmin, rmax;
    rate_range(c, curMax, rmin, rmax);
    double reachHi = curMax * (1 - loss) + rmax;
    double reachLo = curMax * (1 - loss) + rmin;
    if (reachLo <= nextHi && nextLo <= reachHi) return curMax;
    if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
        double upInv = c.x[c.n - 1], upAfter = reachLo;
        for (int i = c.n - 2; i >= 0; i--) {
            double loInv = c.x[i];
            double loAfter = loInv * (1 - loss) + c.lo[i];
            if (loAfter <= nextHi && nextHi <= upAfter) return solve_on_line(loInv, loAfter, upInv, upAfter, nextHi);
            upAfter = loAfter; upInv = loInv;
        }
        throw Error(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    }
    bool found = false; double best = 0.0;
    for (int i = 0; i + 1 < c.n; i++) {
        double rate = c.lo[i];
        double loAfter = c.x[i] * (1 - loss) + rate, upAfter = c.x[i + 1] * (1 - loss) + rate;
        if (loAfter <= nextHi && nextHi <= upAfter) { best = solve_on_line(c.x[i], loAfter, c.x[i + 1], upAfter, nextHi); found = true; }
    }
    if (!found) throw Error(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    return best;
}

double lower_bound_of_space(const Pillars& c, double nextLo, double nextHi, double curMin, double loss) {
    if (c.kind == LSMC_CONSTRAINT_CONSTANT) return std::max((nextLo - c.hi[0]) / (1 - loss), curMin);
    double rmin, rmax;
    rate_range(c, curMin, rmin, rmax);
    double reachHi = curMin * (1 - loss) + rmax;
    double reachLo = curMin * (1 - loss) + rmin;
    if (reachLo <= nextHi && nextLo <= reachHi) return curMin;
    if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
        double loInv = c.x[0], loAfter = reachHi;
        for (int i = 1; i < c.n; i++) {
            double upInv = c.x[i];
            double upAfter = upInv * (1 - loss) + c.hi[i];
            if (loAfter <= nextLo && nextLo <= upAfter) return solve_on_line(loInv, loAfter, upInv, upAfter, nextLo);
            loAfter = upAfter; loInv = upInv;
        }
        throw Error(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    }
    bool found = false; double best = 0.0;
    for (int i = c.n - 2; i >= 0; i--) {
        double rate = c.hi[i];
        double loAfter = c.x[i] * (1 - loss) + rate, upAfter = c.x[i + 1] * (1 - loss) + rate;
        if (loAfter <= nextLo && nextLo <= upAfter) { best = solve_on_line(c.x[i], loAfter, c.x[i + 1], upAfter, nextLo); found = true; }
    }
    if (!found) throw Error(LSMC_ECONSTRAINT, "Storage inventory constraints cannot be satisfied.");
    return best;
}

}  // namespace

void inject_withdraw_range(const lsmc_problem& pb, int period, double inventory, double& rmin, double& rmax) {
    // a breach of a few ulps (rounding of inventory + volume - loss) is accepted, see range_slack()
    if (inventory < pb.min_inventory[period] - range_slack(pb.numerical_tolerance, pb.min_inventory[period]))
        throw Error(LSMC_EINVAL, "Inventory is below minimum allowed value.");
    if (inventory > pb.max_inventory[period] + range_slack(pb.numerical_tolerance, pb.max_inventory[period]))
        throw Error(LSMC_EINVAL, "Inventory above maximum allowed value.");
    if (period >= pb.n_periods - 1) { rmin = 0.0; rmax = 0.0; return; }
    rate_range(pillars_of(pb, period), inventory, rmin, rmax);
}

double decision_cost_npv(const lsmc_problem& pb, int period, double volume) {
    if (volume > 0.0) return (pb.inject_cost[period] * volume) * pb.inject_cost_df[period];
    return (pb.withdraw_cost[period] * std::fabs(-volume)) * pb.withdraw_cost_df[period];
}

double decision_consumed(const lsmc_problem& pb, int period, double volume) {
    if (volume > 0.0) return pb.consumed_inject[period] * std::fabs(volume);
    return pb.consumed_withdraw[period] * std::fabs(-volume);
}

// StorageHelper.CalculateBangBangDecisionSet (SH.cs:109-197).
int bang_bang_decisions(double rmin, double rmax, double inventory, double loss, double next_min, double next_max,
                        double tol, int extra, double* out) {
    if (next_min > next_max) throw Error(LSMC_EINVAL, "nextStepMinInventory cannot be higher than nextStepMaxInventory.");
    const double base = inventory - loss;
    double withdraw, inject;
    const double afterW = rmin + base;
    if (afterW > next_max) {
        if (!(afterW - next_max < tol))
            throw Error(LSMC_EINVAL, "Inventory constraints cannot be fulfilled. This could potentially be fixed by increasing the numerical tolerance.");
        withdraw = next_max - base;
    } else {
        withdraw = afterW > next_min ? rmin : next_min - base;
    }
    const double afterJ = rmax + base;
    if (afterJ < next_min) {
        if (!(next_min - afterJ < tol))
            throw Error(LSMC_EINVAL, "Inventory constraints cannot be fulfilled. This could potentially be fixed by increasing the numerical tolerance.");
        inject = next_min - base;
    } else {
        inject = afterJ < next_max ? rmax : next_max - base;
    }
    auto fill = [](double lo, double hi, int n, double* dst) {
        double step = (hi - lo) / (n + 1);
        for (int i = 0; i < n; i++) dst[i] = (i + 1) * step + lo;
    };
    int count;
    if (withdraw >= 0.0 || inject <= 0.0) {
        count = extra + 2;
        out[0] = withdraw;
        if (extra > 0) fill(withdraw, inject, extra, out + 1);
        out[count - 1] = inject;
    } else {
        count = 2 * extra + 3;
        out[0] = withdraw;
        if (extra > 0) { fill(withdraw, 0, extra, out + 1); fill(0, inject, extra, out + extra + 2); }
        out[extra + 1] = 0.0;
        out[count - 1] = inject;
    }
    return count;
}

// StorageHelper.BisectInventorySpace (SH.cs:280-314).
void bisect_grid(const double* grid, int n, double inventory, double tol, int& lo, int& hi) {
    auto close = [tol](double a, double b) { return std::fabs(a - b) <= tol; };
    if (n == 1 && close(inventory, grid[0])) { lo = hi = 0; return; }
    int left = 0, right = n - 1;
    const int top = right;
    while (right > left) {
        int mid = (left + right) / 2;
        double gm = grid[mid];
        if (close(inventory, gm)) { lo = hi = mid; return; }
        if (gm > inventory) { right = mid; continue; }
        double gn = grid[mid + 1];
        if (inventory <= gn) { lo = mid; hi = mid + 1; return; }
        if (close(inventory, gn)) { lo = hi = mid + 1; return; }
        if (mid + 1 == top) break;
        left = mid;
    }
    throw Error(LSMC_EINVAL, "Inventory is outside of inventoryGrid bounds.");
}

namespace {

// StorageHelper.CalculateInventorySpace (SH.cs:39-107); out[p] for p = 1..P-1, out[0] = start inventory.
void inventory_space(const lsmc_problem& pb, std::vector<double>& lo, std::vector<double>& hi) {
    const int P = pb.n_periods, M = P - 1;
    std::vector<double> fLo(M), fHi(M), bLo(M), bHi(M);
    double runLo = pb.start_inventory, runHi = pb.start_inventory;
    for (int i = 0; i < M; i++) {
        double lossPct = pb.inventory_loss[i], rmin, rmax;
        inject_withdraw_range(pb, i, runLo, rmin, rmax);
        double lossLo = lossPct * runLo;
        runLo = std::max(runLo - lossLo + rmin, pb.min_inventory[i + 1]);
        fLo[i] = runLo;
        inject_withdraw_range(pb, i, runHi, rmin, rmax);
        double lossHi = lossPct * runHi;
        runHi = std::min(runHi - lossHi + rmax, pb.max_inventory[i + 1]);
        fHi[i] = runHi;
    }
    bHi[M - 1] = pb.must_be_empty_at_end ? 0 : pb.max_inventory[P - 1];
    bLo[M - 1] = pb.must_be_empty_at_end ? 0 : pb.min_inventory[P - 1];
    for (int i = M - 2; i >= 0; i--) {
        int period = i + 1;
        Pillars c = pillars_of(pb, period);
        double loss = pb.inventory_loss[period];
        if (c.kind == LSMC_CONSTRAINT_POLYNOMIAL) {
            bHi[i] = polynomial_upper_bound(c, bLo[i + 1], bHi[i + 1], pb.min_inventory[period], pb.max_inventory[period], loss);
            bLo[i] = polynomial_lower_bound(c, bLo[i + 1], bHi[i + 1], pb.min_inventory[period], pb.max_inventory[period], loss);
            continue;
        }
        bHi[i] = upper_bound_of_space(c, bLo[i + 1], bHi[i + 1], pb.max_inventory[period], loss);
        bLo[i] = lower_bound_of_space(c, bLo[i + 1], bHi[i + 1], pb.min_inventory[period], loss);
    }
    lo.assign(P, pb.start_inventory);
    hi.assign(P, pb.start_inventory);
    for (int i = 0; i < M; i++) {
        double top = std::min(fHi[i], bHi[i]), bottom = std::max(fLo[i], bLo[i]);
        if (bottom > top) throw Error(LSMC_ECONSTRAINT, "Inventory constraints cannot be fulfilled.");
        lo[i + 1] = bottom; hi[i + 1] = top;
    }
}

// FixedSpacingStateSpaceGridCalc.GetGridPoints (:45-62): running sum, clamped, one point past the bound.
void append_grid(double lo, double hi, double spacing, std::vector<double>& out) {
    out.push_back(lo);
    if (!(lo < hi)) return;
    double point = lo;
    do {
        point += spacing;
        out.push_back(std::min(point, hi));
    } while (point < hi);
}

inline double expint(double rate, double dt) { return rate == 0.0 ? dt : -std::expm1(-rate * dt) / rate; }

void sim_coefficients(const lsmc_problem& pb, int s0, SimCoefficients& sc) {
    const int F = pb.n_factors, S = pb.n_periods - s0;
    sc.S = S;
    sc.decay.assign((size_t)S * F, 0.0); sc.chol.assign((size_t)S * F * F, 0.0); sc.vol.assign((size_t)S * F, 0.0);
    sc.drift.assign(S, 0.0); sc.fwd.assign(S, 0.0);
    double prev = 0.0;
    std::vector<double> cov((size_t)F * F);
    for (int j = 0; j < S; j++) {
        const int p = j + s0;
        const double t = pb.sim_time[p], dt = t - prev;
        if (!(dt > 0.0)) throw Error(LSMC_EINVAL, "sim_time must be strictly increasing and positive");
        prev = t;
        for (int a = 0; a < F; a++) {
            sc.decay[(size_t)j * F + a] = std::exp(-pb.mean_reversion[a] * dt);
            sc.vol[(size_t)j * F + a] = pb.factor_vol[(size_t)p * F + a];
            for (int b = 0; b < F; b++) cov[a * F + b] = pb.factor_corr[a * F + b] * expint(pb.mean_reversion[a] + pb.mean_reversion[b], dt);
        }
        double* L = &sc.chol[(size_t)j * F * F];
        for (int a = 0; a < F; a++)
            for (int b = 0; b <= a; b++) {
                double s = cov[a * F + b];
                for (int k = 0; k < b; k++) s -= L[a * F + k] * L[b * F + k];
                if (a == b) {
                    if (!(s > 0.0)) throw Error(LSMC_EINVAL, "factor covariance is not positive definite");
                    L[a * F + a] = std::sqrt(s);
                } else L[a * F + b] = s / L[b * F + b];
            }
        double var = 0.0;
        for (int a = 0; a < F; a++)
            for (int b = 0; b < F; b++)
                var += pb.factor_vol[(size_t)p * F + a] * pb.factor_vol[(size_t)p * F + b] * pb.factor_corr[a * F + b] *
                       expint(pb.mean_reversion[a] + pb.mean_reversion[b], t);
        sc.drift[j] = 0.5 * var;
        sc.fwd[j] = pb.forward[p];
    }
}

// Basis monomials as chains over the variable sequence (spot repeated, then each factor repeated); shared
// prefixes become shared nodes, so a full polynomial basis costs about one multiply per monomial.
void compile_basis(const lsmc_problem& pb, BasisProgram& prog) {
    const int F = pb.n_factors, K = pb.n_basis;
    prog.ops.clear();
    prog.node_of_basis.assign(K, 1);
    prog.n_nodes = 3 + F;
    std::map<std::vector<int16_t>, int16_t> known;   // variable sequence -> node
    for (int k = 0; k < K; k++) {
        std::vector<int16_t> seq;
        for (int i = 0; i < pb.spot_power[k]; i++) seq.push_back(2);
        for (int f = 0; f < F; f++)
            for (int i = 0; i < pb.factor_power[k * F + f]; i++) seq.push_back((int16_t)(3 + f));
        if (seq.empty()) { prog.node_of_basis[k] = 1; continue; }
        std::vector<int16_t> prefix{seq[0]};
        int16_t node = seq[0];
        for (size_t i = 1; i < seq.size(); i++) {
            prefix.push_back(seq[i]);
            auto it = known.find(prefix);
            if (it == known.end()) {
                int16_t dst = (int16_t)prog.n_nodes++;
                prog.ops.push_back(BasisOp{dst, node, seq[i], 0});
                known[prefix] = dst;
                node = dst;
            } else node = it->second;
        }
        prog.node_of_basis[k] = node;
    }
}

}  // namespace

void compile_row_program(const BasisProgram& prog, int K, int F, RowProgram& rows) {
    struct Loc { int8_t region; int16_t row; };
    const int n_base = 3 + F;
    std::vector<Loc> loc(prog.n_nodes);
    std::vector<int> level(prog.n_nodes, 0);
    loc[0] = {kRowConst, 0};
    loc[1] = {kRowConst, 1};
    for (int v = 2; v < n_base; v++) loc[v] = {kRowRaw, (int16_t)(v - 2)};
    rows.n_scratch = 0;
    for (int node = n_base; node < prog.n_nodes; node++) {
        int k_first = -1;
        for (int k = 0; k < K && k_first < 0; k++) if (prog.node_of_basis[k] == node) k_first = k;
        if (k_first >= 0) loc[node] = {kRowBasis, (int16_t)k_first};
        else loc[node] = {kRowScratch, (int16_t)rows.n_scratch++};
    }
    struct Pending { RowOp op; int level; };
    std::vector<Pending> pending;
    for (const BasisOp& o : prog.ops) {   // operands precede results in prog.ops (compile_basis)
        level[o.dst] = 1 + std::max(level[o.a], level[o.b]);
        pending.push_back({RowOp{loc[o.a].region, loc[o.b].region, loc[o.dst].region, 0, loc[o.a].row, loc[o.b].row, loc[o.dst].row, 0}, level[o.dst]});
    }
    for (int k = 0; k < K; k++) {
        const int node = prog.node_of_basis[k];
        if (loc[node].region == kRowBasis && loc[node].row == k) continue;   // produced in place
        pending.push_back({RowOp{loc[node].region, kRowConst, kRowBasis, 0, loc[node].row, 1, (int16_t)k, 0}, level[node] + 1});
    }
    std::stable_sort(pending.begin(), pending.end(), [](const Pending& x, const Pending& y) { return x.level < y.level; });
    rows.ops.clear();
    rows.wave_end.clear();
    for (size_t i = 0; i < pending.size(); i++) {
        if (i > 0 && pending[i].level != pending[i - 1].level) rows.wave_end.push_back((int16_t)i);
        rows.ops.push_back(pending[i].op);
    }
    if (!pending.empty()) rows.wave_end.push_back((int16_t)pending.size());
}

void build_plan(const lsmc_problem& problem, Plan& plan) {
    plan.problem.assign(problem);
    const lsmc_problem& pb = plan.problem.pb;
    const int P = pb.n_periods;
    plan.P = P; plan.F = pb.n_factors; plan.K = pb.n_basis;
    plan.s0 = pb.first_period_is_current ? 1 : 0;
    plan.S = P - plan.s0;
    plan.max_dec = 2 * pb.extra_decisions + 3;

    inventory_space(pb, plan.space_min, plan.space_max);

    // grids: period 0 is the single start inventory (LSV.cs:210-211)
    plan.grid.clear(); plan.grid_offset.assign(P + 1, 0);
    plan.max_grid = 1;
    for (int p = 0; p < P; p++) {
        plan.grid_offset[p] = (int32_t)plan.grid.size();
        if (p == 0) plan.grid.push_back(pb.start_inventory);
        else append_grid(plan.space_min[p], plan.space_max[p], pb.grid_spacing, plan.grid);
        plan.max_grid = std::max(plan.max_grid, (int)plan.grid.size() - plan.grid_offset[p]);
    }
    plan.grid_offset[P] = (int32_t)plan.grid.size();

    // decision tables per (period, grid point) (LSV.cs:234-300)
    DecisionTables& tb = plan.tables;
    const int D = plan.max_dec;
    tb.row_offset.assign(P, 0);
    size_t rows = 0;
    for (int p = 0; p < P - 1; p++) { tb.row_offset[p] = (int32_t)rows; rows += plan.grid_size(p); }
    tb.row_offset[P - 1] = (int32_t)rows;
    tb.n_dec.assign(rows, 0); tb.inv_cost_npv.assign(rows, 0.0);
    tb.vol.assign(rows * D, 0.0); tb.cost_npv.assign(rows * D, 0.0); tb.consumed.assign(rows * D, 0.0);
    tb.w_lo.assign(rows * D, 1.0); tb.w_hi.assign(rows * D, 0.0);
    tb.lo.assign(rows * D, 0); tb.hi.assign(rows * D, 0);
    std::vector<double> decisions(D);
    for (int p = 0; p < P - 1; p++) {
        const double* grid = plan.grid_of(p);
        const double* next = plan.grid_of(p + 1);
        const int G = plan.grid_size(p), Gn = plan.grid_size(p + 1);
        const double nextLo = plan.space_min[p + 1], nextHi = plan.space_max[p + 1];
        for (int i = 0; i < G; i++) {
            const size_t row = (size_t)tb.row_offset[p] + i;
            const double v = grid[i];
            double rmin, rmax;
            inject_withdraw_range(pb, p, v, rmin, rmax);
            const double loss = pb.inventory_loss[p] * v;
            const int nd = bang_bang_decisions(rmin, rmax, v, loss, nextLo, nextHi, pb.numerical_tolerance, pb.extra_decisions, decisions.data());
            tb.n_dec[row] = nd;
            tb.inv_cost_npv[row] = pb.inventory_cost[p] == 0.0 ? 0.0 : (v * pb.inventory_cost[p]) * pb.inventory_cost_df[p];
            for (int d = 0; d < nd; d++) {
                const size_t e = row * D + d;
                const double dv = decisions[d];
                tb.vol[e] = dv;
                tb.cost_npv[e] = decision_cost_npv(pb, p, dv);
                tb.consumed[e] = decision_consumed(pb, p, dv);
                const double after = v + dv - loss;
                // linear scan with the hard-coded 1e-8 of LSV.cs:261-299
                bool placed = false;
                // every point more than `margin` below `after` fails both tests of the scan, so it may start just before
                // the first point that is not (same result as starting at 0; the grids are ascending)
                const double margin = 1E-6 + std::fabs(after) * 1E-12;
                int g_start = (int)(std::lower_bound(next, next + Gn, after - margin) - next) - 1;
                if (g_start < 0) g_start = 0;
                for (int g = g_start; g < Gn && !placed; g++) {
                    if (std::fabs(next[g] - after) < 1E-8) {
                        tb.lo[e] = tb.hi[e] = g; tb.w_lo[e] = 1.0; tb.w_hi[e] = 0.0; placed = true;
                    } else if (next[g] > after) {
                        if (g == 0) throw Error(LSMC_EINVAL, "Inventory after decision is below the next period's grid.");
                        double wl = (next[g] - after) / (next[g] - next[g - 1]);
                        tb.lo[e] = g - 1; tb.hi[e] = g; tb.w_lo[e] = wl; tb.w_hi[e] = 1.0 - wl; placed = true;
                    }
                }
                if (!placed) throw Error(LSMC_EINVAL, "Inventory after decision is above the next period's grid.");
            }
        }
    }

    sim_coefficients(pb, plan.s0, plan.sim);
    compile_basis(pb, plan.basis);
    compile_row_program(plan.basis, pb.n_basis, pb.n_factors, plan.basis_rows);
}

// ------------------------------------------------------------------------------------------------
// Intrinsic valuation (IntrinsicStorageValuation.cs:120-322)
// ------------------------------------------------------------------------------------------------
namespace {

// MathNet LinearSpline.Interpolate(x, y).Interpolate(t) as LinearInterpolatorFactory.cs:44-50 uses it: the segment left of t
// (clamped to the first / last segment: linear extrapolation), value y_k + (t - x_k) * slope_k; one point: a constant.
struct LinearInterp {
    const double* x = nullptr;
    std::vector<double> y;
    int n = 0;
    double operator()(double t) const {
        if (n == 1) return y[0];
        int k = (int)(std::upper_bound(x, x + n, t) - x) - 1;
        k = std::min(std::max(k, 0), n - 2);
        const double slope = (y[k + 1] - y[k]) / (x[k + 1] - x[k]);
        return y[k] + (t - x[k]) * slope;
    }
};

struct IntrinsicDecision { double npv, volume, consumed, loss, period_pv; };

// OptimalDecisionAndValue + StorageValueForDecision (IntrinsicStorageValuation.cs:262-320).
template <typename Continuation>
IntrinsicDecision optimal_decision(const lsmc_problem& pb, int p, double inventory, double next_min, double next_max,
                                   double price, const Continuation& continuation) {
    double rmin, rmax;
    inject_withdraw_range(pb, p, inventory, rmin, rmax);
    const double loss = pb.inventory_loss[p] * inventory;
    double ds[3];
    const int nd = bang_bang_decisions(rmin, rmax, inventory, loss, next_min, next_max, pb.numerical_tolerance, 0, ds);
    const double df = pb.settle_df[p];
    IntrinsicDecision best{};
    for (int j = 0; j < nd; j++) {
        const double d = ds[j];
        const double after = inventory + d - loss;
        const double cont = continuation(after);
        const double inject_withdraw_npv = -d * price * df;
        const double inventory_cost_npv = pb.inventory_cost[p] == 0.0 ? 0.0 : (inventory * pb.inventory_cost[p]) * pb.inventory_cost_df[p];
        const double cost_npv = decision_cost_npv(pb, p, d);
        const double consumed = d > 0.0 ? pb.consumed_inject[p] * std::fabs(d) : pb.consumed_withdraw[p] * std::fabs(d);
        const double consumed_npv = -consumed * price * df;
        const double period_pv = inject_withdraw_npv - cost_npv + consumed_npv - inventory_cost_npv;
        const double npv = cont + period_pv;
        if (j == 0 || npv > best.npv) best = IntrinsicDecision{npv, d, consumed, loss, period_pv};   // first strictly greatest (SH.cs:206-222)
    }
    return best;
}

}  // namespace

double intrinsic_value(const Plan& plan, double* profile) {
    const lsmc_problem& pb = plan.problem.pb;
    const int P = plan.P;
    auto terminal = [&](double price, double inv) {
        if (pb.must_be_empty_at_end) return 0.0;
        return pb.terminal_coef[0] + pb.terminal_coef[1] * price + pb.terminal_coef[2] * inv + pb.terminal_coef[3] * (price * inv);
    };
    const double price_end = pb.forward[P - 1];
    // value function of period p (p = 1 .. P-2) on its grid; period P-1 is the terminal value itself
    std::vector<LinearInterp> value(P);
    auto continuation_of = [&](int p_next) {
        return [&, p_next](double inv) { return p_next == P - 1 ? terminal(price_end, inv) : value[p_next](inv); };
    };
    for (int p = P - 2; p >= 1; p--) {
        LinearInterp& v = value[p];
        v.x = plan.grid_of(p); v.n = plan.grid_size(p); v.y.resize(v.n);
        const auto cont = continuation_of(p + 1);
        for (int i = 0; i < v.n; i++)
            v.y[i] = optimal_decision(pb, p, v.x[i], plan.space_min[p + 1], plan.space_max[p + 1], pb.forward[p], cont).npv;
    }
    double inventory = pb.start_inventory, npv = 0.0;
    for (int p = 0; p < P; p++) {
        double* row = profile ? profile + (size_t)p * 6 : nullptr;
        if (p == P - 1) {
            const double end_npv = terminal(pb.forward[p], inventory);
            if (row) { row[0] = inventory; row[1] = row[2] = row[3] = row[4] = 0.0; row[5] = end_npv; }
            npv += end_npv;
        } else {
            const IntrinsicDecision d = optimal_decision(pb, p, inventory, plan.space_min[p + 1], plan.space_max[p + 1], pb.forward[p], continuation_of(p + 1));
            inventory += d.volume - d.loss;
            if (row) { row[0] = inventory; row[1] = d.volume; row[2] = d.consumed; row[3] = d.loss; row[4] = -d.volume - d.consumed; row[5] = d.period_pv; }
            npv += d.period_pv;
        }
    }
    return npv;
}

}  // namespace lsmc
