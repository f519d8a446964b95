// host_plan.h — host-side set-up of one valuation: everything the reference computes per (period, grid point)
// before touching a simulation path (LSV.cs:86-143,209-223,237-300; SH.cs:39-204).  Product code; the CPU
// oracle has its own independent restatement under oracle/.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/lsmc_b200.h"

namespace lsmc {

struct Error : std::runtime_error {
    int status;
    Error(int s, const std::string& m) : std::runtime_error(m), status(s) {}
};

// One multiplication of the basis "program": node[dst] = node[a] * node[b].
struct BasisOp { int16_t dst, a, b, pad; };

// Basis functions compiled to a prefix-sharing product chain (left-to-right, spot first, factors ascending).
// Node 0 = zero row, node 1 = ones, node 2 = spot, node 3+f = factor f, then product nodes.
struct BasisProgram {
    int n_nodes = 0;
    std::vector<BasisOp> ops;
    std::vector<int16_t> node_of_basis;   // [K]
};

// The same program addressed by shared-memory ROWS, for the general sweep (sweep.cu phase F0): every operand and result of
// a multiply is a row of one of four regions, so that a product node which IS a basis function is written straight into
// its design-matrix row.  Copies (basis functions that are a raw variable, the constant, or a second name of a node) are
// multiplies by the ones row: x * 1.0 is exact.  Ops are sorted into waves of mutually independent multiplies.
enum RowRegion : int8_t { kRowConst = 0 /* row 0 zeros, row 1 ones */, kRowRaw = 1 /* 0 spot, 1 + f factor f */, kRowBasis = 2, kRowScratch = 3 };
struct RowOp { int8_t ra, rb, rd, pad; int16_t a, b, d, pad2; };
struct RowProgram {
    std::vector<RowOp> ops;          // wave after wave
    std::vector<int16_t> wave_end;   // [n_waves] one past the last op of each wave
    int n_scratch = 0;               // rows of product nodes that are not basis functions
};
void compile_row_program(const BasisProgram& prog, int K, int F, RowProgram& rows);

// Decision table entry for one (period, grid point, decision): LSV.cs:249-300.
struct DecisionTables {
    // per (period p < P-1, grid index i): offsets via row_offset[p] + i
    std::vector<int32_t> row_offset;      // [P]  first row of period p (rows = grid points)
    std::vector<int32_t> n_dec;           // [rows]
    std::vector<double>  inv_cost_npv;    // [rows]
    // per (row, d), stride max_dec
    std::vector<double> vol, cost_npv, consumed, w_lo, w_hi;
    std::vector<int32_t> lo, hi;
};

struct SimCoefficients {
    int S = 0;
    std::vector<double> decay;   // [S][F]
    std::vector<double> chol;    // [S][F][F]
    std::vector<double> vol;     // [S][F]
    std::vector<double> drift;   // [S]
    std::vector<double> fwd;     // [S]
};

// Deep copy of the caller's problem so no host pointer outlives the call (ownership rule).
struct ProblemCopy {
    lsmc_problem pb{};
    std::vector<double> sim_time, mean_reversion, factor_vol, factor_corr, forward, min_inventory, max_inventory,
        inject_cost, inject_cost_df, withdraw_cost, withdraw_cost_df, consumed_inject, consumed_withdraw, inventory_loss,
        inventory_cost, inventory_cost_df, settle_df, pillar_inventory, pillar_min_rate, pillar_max_rate;
    std::vector<int32_t> spot_power, factor_power, constraint_id, constraint_kind, constraint_first, constraint_count;
    void assign(const lsmc_problem& src);
};

struct Plan {
    ProblemCopy problem;
    int P = 0, F = 0, K = 0, S = 0, s0 = 0, max_dec = 0, max_grid = 0;
    std::vector<double> space_min, space_max;   // [P]; entry 0 is the start inventory
    std::vector<int32_t> grid_offset;           // [P+1]
    std::vector<double> grid;                   // concatenated grids
    DecisionTables tables;
    SimCoefficients sim;
    BasisProgram basis;
    RowProgram basis_rows;

    int grid_size(int p) const { return grid_offset[p + 1] - grid_offset[p]; }
    const double* grid_of(int p) const { return grid.data() + grid_offset[p]; }
};

// Tolerated breach of an inventory bound: the reference checks with none (CmdtyStorage.cs:88-94) and throws on
// rounding noise of `inventory + volume - loss`; see DESIGN.md §6.
inline double range_slack(double tol, double bound) {
    double a = bound < 0 ? -bound : bound;
    double ulps = 64.0 * 2.220446049250313e-16 * (a > 1.0 ? a : 1.0);
    return tol > ulps ? tol : ulps;
}

void build_plan(const lsmc_problem& problem, Plan& plan);

// Intrinsic valuation on the forward curve (IntrinsicStorageValuation.cs:120-322): deterministic backward induction over the
// plan's inventory grids with linear interpolation of the value function (LinearInterpolatorFactory.cs:37-51), then the
// optimal decisions forward from the start inventory.  profile: [P][6] as StorageProfile (inventory AFTER the period's
// decision, :283-286).  Host only: O(P * G * 3).
double intrinsic_value(const Plan& plan, double* profile);

// Storage queries at arbitrary inventory, used by the trigger-price post-processing (LSV.cs:494-497).
void inject_withdraw_range(const lsmc_problem& pb, int period, double inventory, double& rmin, double& rmax);
int bang_bang_decisions(double rmin, double rmax, double inventory, double loss, double next_min, double next_max,
                        double tol, int extra, double* out);
double decision_cost_npv(const lsmc_problem& pb, int period, double volume);
double decision_consumed(const lsmc_problem& pb, int period, double volume);
void bisect_grid(const double* grid, int n, double inventory, double tol, int& lo, int& hi);

}  // namespace lsmc
