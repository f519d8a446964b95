// lsmc_device.cuh — device helpers and kernel parameter blocks for the sm_100a LSMC kernels.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lsmc {

constexpr int kMaxFactors = 8;
constexpr int kMaxNodes = 200;    // basis-program nodes (zero, one, spot, F factors, products)
constexpr int kMaxOps = 192;

// ------------------------------------------------------------------------------------------------
// Programmatic dependent launch (PDL): the kernels of the backward loop are launched with programmatic stream
// serialisation (kernels.h launch_pdl), so a kernel's CTAs are scheduled while its predecessor drains and its set-up that
// does not touch the predecessor's results (shared-memory staging, decision tables, the first regressor copies) overlaps
// it.  pdl_wait() returns when every prerequisite grid has completed and its memory operations are visible; it is a
// no-op in a kernel launched without the attribute.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------
// Arithmetic with the reference's rounding: no contraction unless written as fma().
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }

// Deterministic exp: the same IEEE operation sequence as the CPU oracle's det_exp, so injected-normal
// simulations agree bit for bit (the reference's Math.Exp is a third-party CRT routine).
__device__ __forceinline__ double det_exp(double x) {
    if (x > 709.0) return __longlong_as_double(0x7ff0000000000000LL);
    if (x < -708.0) return 0.0;
    const double LOG2E = 1.4426950408889634074;
    const double LN2_HI = 6.93147180369123816490e-01;
    const double LN2_LO = 1.90821492927058770002e-10;
    double kf = rint(__dmul_rn(x, LOG2E));
    double r = __fma_rn(-kf, LN2_HI, x);
    r = __fma_rn(-kf, LN2_LO, r);
    double p = 1.6059043836821613e-10;
    p = __fma_rn(p, r, 2.08767569878681e-09);
    p = __fma_rn(p, r, 2.505210838544172e-08);
    p = __fma_rn(p, r, 2.755731922398589e-07);
    p = __fma_rn(p, r, 2.7557319223985893e-06);
    p = __fma_rn(p, r, 2.48015873015873e-05);
    p = __fma_rn(p, r, 0.0001984126984126984);
    p = __fma_rn(p, r, 0.001388888888888889);
    p = __fma_rn(p, r, 0.008333333333333333);
    p = __fma_rn(p, r, 0.041666666666666664);
    p = __fma_rn(p, r, 0.16666666666666666);
    p = __fma_rn(p, r, 0.5);
    p = __fma_rn(p, r, 1.0);
    p = __fma_rn(p, r, 1.0);
    long long k = (long long)kf;
    double scale = __longlong_as_double((k + 1023) << 52);
    return __dmul_rn(p, scale);
}

// Philox4x32-10 (Salmon et al., SC'11).
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
        uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
        uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += W0; k1 += W1;
    }
}

__device__ __forceinline__ void philox_normal_pair(uint64_t seed, uint64_t draw, uint32_t step, uint32_t pair, double& z0, double& z1) {
    uint32_t c[4] = {(uint32_t)draw, (uint32_t)(draw >> 32), step, pair};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    uint64_t a = ((uint64_t)c[0] << 32) | c[1];
    uint64_t b = ((uint64_t)c[2] << 32) | c[3];
    double u1 = ((double)(a >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    double u2 = ((double)(b >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    double rad = sqrt(-2.0 * log(u1));
    double sn, cs;
    sincospi(2.0 * u2, &sn, &cs);
    z0 = rad * cs;
    z1 = rad * sn;
}

// FP64 tensor-core tile: D(8x8) += A(8x4, row) * B(4x8, col).  SASS: DMMA.8x8x4.
// lane = 4*g + t:  A[g][t], B[t][g], D[g][2t], D[g][2t+1].
__device__ __forceinline__ void dmma8x8x4(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

struct BasisOpDev { int16_t dst, a, b, pad; };
struct RowOpDev { int8_t ra, rb, rd, pad; int16_t a, b, d, pad2; };   // host_plan.h RowOp: (region, row) of the operands and the result
constexpr int kMaxRowOps = kMaxOps + 64;

// Compiled basis (host_plan.h BasisProgram) + exponents, uploaded once per problem.
struct BasisDev {
    int n_nodes, n_ops, K, F;
    BasisOpDev ops[kMaxOps];
    int16_t node_of_basis[64];
    int8_t spot_pow[64];                 // exponents (BasisFunctions.cs:92-132), used by the preconditioner kernel
    int8_t fac_pow[64][kMaxFactors];
    // the same program by shared-memory rows, in waves (host_plan.h RowProgram), for the general sweep
    int n_rowops, n_rowwaves, n_scratch;
    int16_t rowwave_end[kMaxRowOps];
    RowOpDev rowops[kMaxRowOps];
    // `ops` in waves of mutually independent multiplies (a wave ends before the first op that reads a result of the same wave;
    // compile_basis emits operands first): used by the pipelined sweep, computed once on the host
    int n_opwaves;
    int16_t opwave_end[kMaxOps + 1];
};

// Per-period regression preconditioner: T = R_sub^-1 (upper triangular, K x K, row-major) from a
// Householder QR of the design matrix on a subsample of paths, followed by K dropped-column flags.
__host__ __device__ inline size_t precond_stride(int K) { return (size_t)K * K + K; }

// ------------------------------------------------------------------------------------------------
// Parameter blocks
// ------------------------------------------------------------------------------------------------
struct SimParams {
    int64_t N, Npad, sim_offset;   // local paths, row stride, global index of path 0
    int S, F, mode, antithetic;
    uint64_t seed;
    const double* normals;               // [S][F][N] (mode NORMALS), device
    const double* decay; const double* chol; const double* vol; const double* drift; const double* fwd;
    double* spot;                        // [S][Npad]
    double* fac;                         // [S][F][Npad]
};

struct SweepParams {
    int64_t N, Npad;
    int G, Gn;                // grid sizes of this period and of the next
    int K, F, max_dec;
    int mode;                 // 0 decide, 1 terminal value
    int cont_mode;            // 0 fitted from beta, 1 per-row constants (current period: means)
    int accum_mode;           // 0 none, 1 basis of period p-1, 2 ones (column sums)
    int spot_is_const;
    double spot_const, df;
    double tc0, tc1, tc2, tc3;   // terminal value coefficients
    int must_be_empty;
    const double* Vnext;      // [Gn][Npad]
    double* Vout;             // [G][Npad]
    const double* spot_p; const double* fac_p;    // period p regressors ([Npad], [F][Npad])
    const double* spot_m; const double* fac_m;    // period p-1 regressors
    const double* beta;       // [Gn][ldb] fitted coefficients (cont_mode 0) or [Gn] constants at stride ldb (cont_mode 1)
    int ldb;
    const double* grid;       // [G] inventories of this period (terminal mode)
    // decision tables for this period (row = grid point, stride max_dec)
    const int32_t* t_ndec; const double* t_invcost;
    const double* t_vol; const double* t_cost; const double* t_cons; const double* t_wlo; const double* t_whi;
    const int32_t* t_lo; const int32_t* t_hi;
    const BasisDev* basis;
    const double* precond;    // preconditioner of period p-1 (accum_mode 1): T [K][K] + flags [K]
    double* partial;          // [gridDim.x][8*MT][ncols] normal-equation partials
    int ncols;                // 8 * (ceil(G/8) + MT)
    long long* trace;         // development (LSMC_TRACE): per-warp phase timestamps of CTA 0, [16][64][8]; NULL otherwise
    int pdl;                  // host side: launch with programmatic stream serialisation (kernels.h launch_pdl)
    int n_prod;               // product nodes of the basis program (= basis->n_ops; a parameter so that the kernel's set-up does not start with a load)
};

struct PrecondParams {
    int64_t N, Npad;
    int S, K, F, nsub;
    const double* spot;       // [S][Npad] regression set
    const double* fac;        // [S][F][Npad]
    const BasisDev* basis;
    double* out;              // [S][precond_stride(K)]
};

struct ForwardParams {
    int64_t N, Npad;
    int p, Gn, K, F, max_dec, extra;
    int cont_mode, spot_is_const;
    double spot_const, df, tol;
    double loss_pct, inv_cost, inv_cost_df, inj_cost, inj_df, wd_cost, wd_df, cons_inj, cons_wd;
    double min_inv, max_inv, next_min, next_max, slack_lo, slack_hi;
    int c_kind, c_n; const double* c_x; const double* c_lo; const double* c_hi;
    const double* spot; const double* fac;   // valuation regressors of period p
    const double* beta; int ldb;
    const double* next_grid;
    const BasisDev* basis;
    double* inventory;        // [Npad] state in/out
    double* pv_by_sim;        // [Npad] accumulated
    double* sums;             // [gridDim.x][kFwdSums + K] block partials
    int* error_flag;
    // optional per-sim panels rows for this period (NULL to skip)
    double* o_inventory; double* o_volume; double* o_consumed; double* o_loss; double* o_net; double* o_pv;
    int8_t* o_decision;
};
constexpr int kFwdSums = 7;   // inventory, volume, consumed, loss, net volume, pv, -(d+mu)*s

}  // namespace lsmc
