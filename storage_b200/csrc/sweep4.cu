// sweep4.cu — K3, pipelined form: the fused backward sweep of one regular period (LSV.cs:185-205, 234-332, 753-770).
//
// Same arithmetic as sweep.cu (the general kernel, kept for the terminal / current-period launches and for shapes this
// one does not cover), reorganised as a warp-specialised software pipeline with no CTA-wide barrier in the tile loop.
// One persistent CTA per SM walks tiles of TP paths; 16 warps:
//   warps 0..12   "decision" warps: warp j owns grid rows 8j..8j+7 for the whole kernel, its decision table lives in
//                 registers.  Lane (g,t) handles row 8j+g and, per iteration, paths 8u+2t, 8u+2t+1: max over the
//                 bang-bang decisions of immediate + interpolated fitted continuation, realised value from the
//                 interpolated V_next, 128-bit store of V, and the value goes straight from registers into the
//                 FP64 tensor-pipe accumulation of X~^T V (DMMA 8x8x4 for whole blocks of 8 basis rows, DFMA for a
//                 remainder of 1..3 rows: on B200 both run on the same 64-lane pipe, so padding 10 rows to 16 would
//                 cost 60% more pipe time for nothing).
//   warps 13..15  "aux" warps, one tile ahead of the decision warps:
//                 warp 13: producer — one TMA tensor copy (cp.async.bulk.tensor.2d) per V_next tile (2 buffers) and
//                          four per raw-regressor stage (spot/factors of periods p and p-1, 3-stage ring); then the
//                          preconditioned rows x~ = T^T x_{p-1} (DMMA), X~^T X~ (DMMA) and one block of fitted values;
//                 warps 14, 15: the product nodes of the basis monomials of period p / p-1 (one multiply per node,
//                          host_plan.cpp compile_basis), then the fitted continuation values C^ = beta . x_p on the
//                          FP64 tensor pipe with the beta fragments register resident, into the other C^ buffer.
// Hand-offs are mbarriers: vfull/rfull (copies landed), cready (aux -> decision), dfree (decision -> aux).
// The warps of one SM sub-partition are 4 decision warps, or 3 decision warps + 1 aux warp: the FP64 pipe — which,
// not HBM, is the tighter bound of this kernel — carries about the same load on all four.
#include <cuda.h>
#include <type_traits>

#include "kernels.h"

namespace lsmc {

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(addr), "r"(parity) : "memory");
}
// 2-D tiled TMA copy global -> shared: box (TP paths x rows) at element coordinates (c0 = path, c1 = row).
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void aux_bar() { asm volatile("bar.sync 1, 96;" ::: "memory"); }
__device__ __forceinline__ double2 lds128(const double* p) { return *reinterpret_cast<const double2*>(p); }

// Unroll factor of the 8-path step loop of a decision warp's tile.  1 (rolled) is the measured optimum: with the whole
// tile straight-line (7) ptxas sinks the seven V stores to the end of the block and holds their 28 registers, spills the
// row pointers, and the launch is 9 % slower (0.570 -> 0.533 ms at 1M paths; 2: 0.548, 3: 0.580).
#ifndef LSMC_STEP_UNROLL
#define LSMC_STEP_UNROLL 1
#endif
constexpr int kStepUnroll = LSMC_STEP_UNROLL;
// (Weight and row offsets of the chosen decision come from register selects; a per-warp selection table in shared memory,
// software-pipelined and register-prefetched step loops were measured slower — DESIGN.md 4.1 — and are gone from the source.)
#ifndef LSMC_ABLATE   // development only: timing ablations (results are wrong when != 0)
#define LSMC_ABLATE 0
#endif
// 5: no decisions; 6: no aux DMMA work; 7: 5 + no fit; 8: 5 + no x~ / Gram; 9: 5 + 6 (TMA + barrier skeleton only)
#define AB_NO_DECIDE (LSMC_ABLATE == 5 || LSMC_ABLATE == 7 || LSMC_ABLATE == 8 || LSMC_ABLATE == 9)
#define AB_NO_XT (LSMC_ABLATE == 6 || LSMC_ABLATE == 8 || LSMC_ABLATE == 9)
#define AB_NO_FIT (LSMC_ABLATE == 6 || LSMC_ABLATE == 7 || LSMC_ABLATE == 9)
// 13: only decision warps 0..6 decide; 14: only warps 0, 4, 8, 12 (one sub-partition) decide

constexpr int kWarps = 16;
constexpr int kDecWarps = 13;          // grid rows <= 104
constexpr int kAuxWarps = 3;
constexpr int kRawStages = 3;
constexpr int kMaxProd = 32;           // product nodes of the basis program
constexpr int kElemReg = 8;          // product elements per aux lane kept as register descriptors
constexpr int kFitTiles = 6;           // C^ row blocks per fit warp (warps 14, 15; warp 13 takes the 13th)

__host__ __device__ inline size_t al128(size_t b) { return (b + 127) & ~(size_t)127; }

// Shared-memory carve-up in BYTES; every TMA destination is 128-byte aligned.
struct Layout4 {
    int GnP, MT, LDP, VP;                                 // VP: pitch of the V_next tile in paths (= TP: rows 448 bytes apart alternate bank halves, 128-bit gathers conflict free)
    size_t o_sp, o_fp, o_sm, o_fm, stage;                 // raw stage: spot_p | fac_p[F] | spot_m | fac_m[F]
    size_t vn, vn_buf, c, raw, prod, xt, zero, ones, ones_last, total;   // vn_buf: bytes of one V_next buffer
};

__host__ __device__ inline Layout4 layout4(int TP, int Gn, int K, int F, int n_prod) {
    Layout4 L;
    L.GnP = (Gn + 7) & ~7;
    L.MT = (K + 7) >> 3;
    L.LDP = TP + 4;
    L.VP = TP;
    const size_t row = (size_t)TP * 8;
    L.o_sp = 0; L.o_fp = al128(row); L.o_sm = L.o_fp + al128(F * row); L.o_fm = L.o_sm + al128(row);
    L.stage = L.o_fm + al128(F * row);
    size_t o = 0;
    L.vn_buf = al128((size_t)Gn * L.VP * 8);
    L.vn = o;   o += 2 * L.vn_buf;
    L.c = o;    o += al128((size_t)2 * L.GnP * row);
    L.raw = o;  o += kRawStages * L.stage;
    L.prod = o; o += al128((size_t)2 * (n_prod > 0 ? n_prod : 1) * L.LDP * 8);
    L.xt = o;   o += al128((size_t)2 * 8 * L.MT * row);
    L.zero = o; o += al128(row);
    L.ones = o; o += al128(row);
    L.ones_last = o; o += al128(row);
    L.total = o;
    return L;
}

// Row descriptor of a basis-program node: byte offset, bit 31 = relative to the current raw stage, bit 30 = ones row
// (replaced by the masked ones row in the tile that contains path N).
constexpr uint32_t kRel = 0x80000000u, kOnes = 0x40000000u, kOffMask = 0x3fffffffu;

struct RowTab {
    int row_ok, variant;
    double invc;
    double nvol[3], cost[3], ncons[3], wl[3], wh[3];   // nvol = -volume, ncons = -consumed (exact negations)
    int rL[3], rH[3];                                  // interpolation rows of the next grid
};

enum : int { kZero1 = 1, kCons0 = 2, kCons2 = 4, kInvc = 8, kGeneric = 16 };

// Shared state of one CTA's pipeline, handed to the per-variant decision loops.
struct Pipe4 {
    unsigned char* sm;
    uint64_t *vfull, *cready, *dfree, *rfull;
};

// Timing ablations: a decision warp that only keeps the pipeline's hand-shakes going.
__device__ __noinline__ void decision_loop_empty(const SweepParams& q, const Pipe4& P, int TP) {
    const int64_t n_tiles = (q.N + TP - 1) / TP;
    int stage = 0, sphase = 0, it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
        const int b = it & 1;
        const uint32_t par = (uint32_t)((it >> 1) & 1);
        mbar_wait(&P.cready[b], par);
        mbar_wait(&P.rfull[stage], (uint32_t)sphase);
        mbar_wait(&P.vfull[b], par);
        mbar_arrive(&P.dfree[b]);
        if (++stage == kRawStages) { stage = 0; sphase ^= 1; }
    }
}

// The decision warps' tile loop for one table variant.  tot_d = ((-vol_d s) df - cost_d [+ (-cons_d s) df]) + cont_d [- invc]
// (LSV.cs:304-330, operation order kept; the skipped terms are exact no-ops, see sweep.cu).  Each 8-path step is split
// into phase A (fitted values -> totals -> first strictly greatest, SH.cs:206-222) and phase B (weights and rows of the
// chosen decision from the warp's table in shared memory, gather of V_next, value, store, normal-equation update); phase B
// of step u-1 is issued after phase A of step u, so the two dependent shared-memory round trips of the gather hide
// behind independent FP64 work.
template <int TP, int MTD, int KS, int FLAGS>
__device__ __forceinline__ void decision_loop(const SweepParams& q, const Layout4& L, const Pipe4& P, const RowTab& R,
                                              const int j, const int g, const int t) {
    constexpr bool ZERO1 = (FLAGS & kZero1) != 0, CONS0 = (FLAGS & kCons0) != 0, CONS2 = (FLAGS & kCons2) != 0, INVC = (FLAGS & kInvc) != 0;
    constexpr bool CONS1 = CONS0 || CONS2;
    constexpr int MT = MTD + (KS > 0 ? 1 : 0);
    constexpr int NIT = TP / 8;
    const int GnP = L.GnP;
    const int64_t n_tiles = (q.N + TP - 1) / TP;
    const double df = q.df;
    double acc[MTD > 0 ? MTD : 1][2];
    double accS[KS > 0 ? KS : 1];
#pragma unroll
    for (int m = 0; m < (MTD > 0 ? MTD : 1); m++) acc[m][0] = acc[m][1] = 0.0;
#pragma unroll
    for (int r = 0; r < (KS > 0 ? KS : 1); r++) accS[r] = 0.0;
    const int i = 8 * j + g;
    double* const vrow = q.Vout + (size_t)(R.row_ok ? i : 0) * q.Npad + 2 * t;

    // byte offsets of the interpolation rows of each decision in the V_next tile (pitch TP), lo | hi << 16
    const uint32_t voff0 = (uint32_t)(R.rL[0] * TP * 8) | ((uint32_t)(R.rH[0] * TP * 8) << 16);
    const uint32_t voff1 = (uint32_t)(R.rL[1] * TP * 8) | ((uint32_t)(R.rH[1] * TP * 8) << 16);
    const uint32_t voff2 = (uint32_t)(R.rL[2] * TP * 8) | ((uint32_t)(R.rH[2] * TP * 8) << 16);
    int stage = 0, sphase = 0, it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
        const int b = it & 1;
        const uint32_t par = (uint32_t)((it >> 1) & 1);
        // development (LSMC_TRACE): phase timestamps of CTA 0, read by scripts/trace_report.py
        long long* tr = (q.trace && blockIdx.x == 0 && it < 64 && (threadIdx.x & 31) == 0) ? q.trace + ((size_t)(threadIdx.x >> 5) * 64 + it) * 8 : nullptr;
        if (tr) tr[0] = clock64();
        mbar_wait(&P.cready[b], par);                   // C^ and x~ of this tile are in place
        if (tr) tr[1] = clock64();
        mbar_wait(&P.rfull[stage], (uint32_t)sphase);   // spot row (landed long ago; makes the async-proxy writes visible here)
        mbar_wait(&P.vfull[b], par);                    // V_next tile
        if (tr) tr[2] = clock64();
        const double* sC = reinterpret_cast<const double*>(P.sm + L.c) + (size_t)b * GnP * TP + 2 * t;
        const unsigned char* pV = P.sm + L.vn + (size_t)b * L.vn_buf + 16 * t;
        const double* pS = reinterpret_cast<const double*>(P.sm + L.raw + (size_t)stage * L.stage + L.o_sp) + 2 * t;
        const double* sX = reinterpret_cast<const double*>(P.sm + L.xt) + (size_t)b * 8 * MT * TP + 2 * t;
        const double* pL0 = sC + R.rL[0] * TP; const double* pH0 = sC + R.rH[0] * TP;
        const double* pL1 = sC + R.rL[1] * TP; const double* pH1 = sC + R.rH[1] * TP;
        const double* pL2 = sC + R.rL[2] * TP; const double* pH2 = sC + R.rH[2] * TP;
        const double* pXd = sX + (size_t)g * TP;            // DMMA A operand: x~[8m+g][.]
        const double* pXs = sX + (size_t)(8 * MTD) * TP;    // scalar remainder rows x~[8 MTD + r][.]
        double* out = vrow + tile * TP;

        double best0, best1, bc0, bc1;   // carried from phase A to phase B
        double sw0, sw1;                 // w_lo of the chosen decision
        uint32_t k0, k1;                 // byte offsets of its interpolation rows in the V_next tile: lo | hi << 16
        struct StepIn { double2 s2, cl[3], ch[3]; };   // spot and the bracketing fitted values of the three decisions, two paths
        auto loadA = [&](const int u8, StepIn& in) {
            in.s2 = lds128(pS + u8);
            in.cl[0] = lds128(pL0 + u8); in.ch[0] = lds128(pH0 + u8);
            in.cl[1] = lds128(pL1 + u8); in.ch[1] = lds128(pH1 + u8);
            in.cl[2] = lds128(pL2 + u8); in.ch[2] = lds128(pH2 + u8);
        };
        auto computeA = [&](const StepIn& in) {
            const double2 s2 = in.s2;
            double c0[3], c1[3], x0[3], x1[3];
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double2 cl = in.cl[d];
                const double2 ch = in.ch[d];
                c0[d] = add_rn(mul_rn(ch.x, R.wh[d]), mul_rn(cl.x, R.wl[d]));
                c1[d] = add_rn(mul_rn(ch.y, R.wh[d]), mul_rn(cl.y, R.wl[d]));
                if (d == 1 && ZERO1) { x0[d] = c0[d]; x1[d] = c1[d]; }
                else {
                    double i0 = sub_rn(mul_rn(mul_rn(R.nvol[d], s2.x), df), R.cost[d]);
                    double i1 = sub_rn(mul_rn(mul_rn(R.nvol[d], s2.y), df), R.cost[d]);
                    if ((d == 0 && CONS0) || (d == 1 && CONS1) || (d == 2 && CONS2)) {
                        i0 = add_rn(i0, mul_rn(mul_rn(R.ncons[d], s2.x), df));
                        i1 = add_rn(i1, mul_rn(mul_rn(R.ncons[d], s2.y), df));
                    }
                    x0[d] = add_rn(i0, c0[d]); x1[d] = add_rn(i1, c1[d]);
                }
                if (INVC) { x0[d] = sub_rn(x0[d], R.invc); x1[d] = sub_rn(x1[d], R.invc); }
            }
            const bool a1 = x0[1] > x0[0], b1 = x1[1] > x1[0];
            best0 = a1 ? x0[1] : x0[0]; best1 = b1 ? x1[1] : x1[0];
            bc0 = a1 ? c0[1] : c0[0]; bc1 = b1 ? c1[1] : c1[0];
            k0 = a1 ? voff1 : voff0; k1 = b1 ? voff1 : voff0;
            sw0 = a1 ? R.wl[1] : R.wl[0]; sw1 = b1 ? R.wl[1] : R.wl[0];
            const bool a2 = x0[2] > best0, b2 = x1[2] > best1;
            best0 = a2 ? x0[2] : best0; best1 = b2 ? x1[2] : best1;
            bc0 = a2 ? c0[2] : bc0; bc1 = b2 ? c1[2] : bc1;
            k0 = a2 ? voff2 : k0; k1 = b2 ? voff2 : k1;
            sw0 = a2 ? R.wl[2] : sw0; sw1 = b2 ? R.wl[2] : sw1;
        };
        auto phaseA = [&](const int u8) { StepIn in; loadA(u8, in); computeA(in); };
        auto phaseB = [&](const int u8, const double bst0, const double bst1, const double bcc0, const double bcc1, const uint32_t kk0, const uint32_t kk1,
                          const double wl0, const double wl1) {
            // 128-bit gathers: the rows of path 0's decision serve path 1 too when both paths chose the same decision
            const double2 gl0 = *reinterpret_cast<const double2*>(pV + (kk0 & 0xffffu) + 8 * u8), gh0 = *reinterpret_cast<const double2*>(pV + (kk0 >> 16) + 8 * u8);
            double2 gl1 = gl0, gh1 = gh0;
            if (kk1 != kk0) {
                gl1 = *reinterpret_cast<const double2*>(pV + (kk1 & 0xffffu) + 8 * u8);
                gh1 = *reinterpret_cast<const double2*>(pV + (kk1 >> 16) + 8 * u8);
            }
            const double vl0 = gl0.x, vh0 = gh0.x, vl1 = gl1.y, vh1 = gh1.y;
            // w_hi = 1 - w_lo exactly as host_plan.cpp build_tables forms it (LSV.cs:277: wu = 1 - wl)
            const double act0 = add_rn(mul_rn(vh0, sub_rn(1.0, wl0)), mul_rn(vl0, wl0));
            const double act1 = add_rn(mul_rn(vh1, sub_rn(1.0, wl1)), mul_rn(vl1, wl1));
            double v0 = add_rn(bst0, add_rn(-bcc0, act0));
            double v1 = add_rn(bst1, add_rn(-bcc1, act1));
            // lanes of rows beyond the grid compute on row 0 (valid memory, finite values) and contribute zeros
            if (R.row_ok) __stcs(reinterpret_cast<double2*>(out + u8), make_double2(v0, v1));
            else { v0 = 0.0; v1 = 0.0; }
#pragma unroll
            for (int m = 0; m < MTD; m++) {
                const double2 a2 = lds128(pXd + (size_t)8 * m * TP + u8);
                dmma8x8x4(acc[m][0], acc[m][1], a2.x, v0);
                dmma8x8x4(acc[m][0], acc[m][1], a2.y, v1);
            }
#pragma unroll
            for (int r = 0; r < KS; r++) {
                const double2 a2 = lds128(pXs + (size_t)r * TP + u8);
                accS[r] = fma(a2.x, v0, accS[r]);
                accS[r] = fma(a2.y, v1, accS[r]);
            }
        };
#pragma unroll (kStepUnroll)
        for (int u = 0; u < NIT; u++) {
            phaseA(8 * u);
            phaseB(8 * u, best0, best1, bc0, bc1, k0, k1, sw0, sw1);
        }
        if (tr) tr[3] = clock64();
        mbar_arrive(&P.dfree[b]);
        if (++stage == kRawStages) { stage = 0; sphase ^= 1; }
    }
    // per-CTA partial: partial[cta][k][col]; DMMA blocks: k = 8m+g, col = 8j+2t; scalar rows: k = 8 MTD + r, col = 8j+g
    double* po = q.partial + (size_t)blockIdx.x * (8 * MT) * q.ncols;
#pragma unroll
    for (int m = 0; m < MTD; m++)
        *reinterpret_cast<double2*>(po + (size_t)(8 * m + g) * q.ncols + 8 * j + 2 * t) = make_double2(acc[m][0], acc[m][1]);
#pragma unroll
    for (int r = 0; r < KS; r++) {
        double v = accS[r];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (t == 0) po[(size_t)(8 * MTD + r) * q.ncols + 8 * j + g] = v;
    }
    if (KS > 0 && t != 0) {   // rows K .. 8 MT - 1 of the scalar block are padding: keep them zero
        for (int r = KS + t - 1; r < 8; r += 3) po[(size_t)(8 * MTD + r) * q.ncols + 8 * j + g] = 0.0;
    }
}

}  // namespace

template <int TP, int MTD, int KS, int KSTEPS>
__global__ void __launch_bounds__(kWarps * 32, 1) sweep4_kernel(const SweepParams q, const __grid_constant__ CUtensorMap mapV,
                                                                const __grid_constant__ CUtensorMap mapSp, const __grid_constant__ CUtensorMap mapFp,
                                                                const __grid_constant__ CUtensorMap mapSm, const __grid_constant__ CUtensorMap mapFm) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // PDL: the kernel after this one (tail.cu) may be scheduled as soon as SMs free up; it waits for this grid's completion itself
    pdl_launch_dependents();
    if (q.trace && blockIdx.x == 0 && threadIdx.x == 0) q.trace[(0 * 64 + 63) * 8 + 6] = clock64();   // kernel entry (development)
    constexpr int MT = MTD + (KS > 0 ? 1 : 0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int K = q.K, F = q.F;
    const BasisDev* __restrict__ bd = q.basis;
    const int n_prod = q.n_prod;
    const Layout4 L = layout4(TP, q.Gn, K, F, n_prod);
    const int GnP = L.GnP;
    constexpr size_t ROW = (size_t)TP * 8;
    unsigned char* const sm = smem_raw;
    __shared__ uint64_t vfull[2], cready[2], dfree[2], rfull[kRawStages];
    __shared__ uint32_t sNode[2][3 + kMaxFactors + kMaxProd];   // row descriptor of every node, per period (0: p, 1: p-1)
    __shared__ uint32_t sOpA[2][kMaxProd], sOpB[2][kMaxProd], sOpD[2][kMaxProd];
    __shared__ int sWaveEnd[kMaxProd + 1], sNWaves;   // product nodes in waves of mutually independent multiplies

    const int NT = (q.G + 7) >> 3;                 // decision warps in use
    const int64_t n_tiles = (q.N + TP - 1) / TP;

    // ---- one-time staging (all threads) ----
    for (size_t i = tid; i < L.total / 8; i += blockDim.x) reinterpret_cast<double*>(sm)[i] = 0.0;
    for (int i = tid; i < 2 * (3 + F + n_prod); i += blockDim.x) {
        const int sel = i / (3 + F + n_prod), id = i - sel * (3 + F + n_prod);
        uint32_t d;
        if (id == 0) d = (uint32_t)L.zero;
        else if (id == 1) d = (uint32_t)L.ones | kOnes;
        else if (id == 2) d = (uint32_t)(sel ? L.o_sm : L.o_sp) | kRel;
        else if (id < 3 + F) d = (uint32_t)((sel ? L.o_fm : L.o_fp) + (size_t)(id - 3) * ROW) | kRel;
        else d = (uint32_t)(L.prod + ((size_t)sel * n_prod + (id - 3 - F)) * L.LDP * 8);
        sNode[sel][id] = d;
    }
    __syncthreads();
    for (int i = tid; i < TP; i += blockDim.x) {
        reinterpret_cast<double*>(sm + L.ones)[i] = 1.0;
        reinterpret_cast<double*>(sm + L.ones_last)[i] = ((n_tiles - 1) * TP + i < q.N) ? 1.0 : 0.0;
    }
    for (int i = tid; i < 2 * n_prod; i += blockDim.x) {
        const int sel = i / n_prod, o = i - sel * n_prod;
        const BasisOpDev op = bd->ops[o];
        sOpA[sel][o] = sNode[sel][op.a]; sOpB[sel][o] = sNode[sel][op.b]; sOpD[sel][o] = sNode[sel][op.dst];
    }
    if (tid == 0) {
        for (int b = 0; b < 2; b++) { mbar_init(&vfull[b], 1); mbar_init(&cready[b], kAuxWarps * 32); mbar_init(&dfree[b], NT * 32); }
        for (int s = 0; s < kRawStages; s++) mbar_init(&rfull[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // waves of mutually independent product nodes: computed once on the host (BasisDev::opwave_end)
    for (int w = tid; w < bd->n_opwaves; w += blockDim.x) sWaveEnd[w] = bd->opwave_end[w];
    if (tid == 32) sNWaves = bd->n_opwaves;
    __syncthreads();

    if (warp < kDecWarps) {
        // =========================================== decision warps ===========================================
        if (warp >= NT) return;
        const int j = warp, i = 8 * j + g;
        RowTab R;
        R.row_ok = (i < q.G) ? 1 : 0;
        {
            const bool on = R.row_ok != 0;
            const size_t trow = (size_t)(on ? i : 0) * q.max_dec;
            const int nd = on ? __ldg(q.t_ndec + i) : 0;
            R.invc = on ? __ldg(q.t_invcost + i) : 0.0;
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const int ds = (d < nd) ? d : 0;   // missing decisions repeat decision 0 (a strict '>' never selects them)
                R.nvol[d] = on ? -__ldg(q.t_vol + trow + ds) : 0.0;
                R.cost[d] = on ? __ldg(q.t_cost + trow + ds) : 0.0;
                R.ncons[d] = on ? -__ldg(q.t_cons + trow + ds) : 0.0;
                R.wl[d] = on ? __ldg(q.t_wlo + trow + ds) : 1.0;
                R.wh[d] = on ? __ldg(q.t_whi + trow + ds) : 0.0;
                R.rL[d] = on ? __ldg(q.t_lo + trow + ds) : 0;
                R.rH[d] = on ? __ldg(q.t_hi + trow + ds) : 0;
            }
            const bool zero1 = __all_sync(0xffffffffu, R.nvol[1] == 0.0 && R.cost[1] == 0.0 && R.ncons[1] == 0.0);
            const bool cons0 = __any_sync(0xffffffffu, R.ncons[0] != 0.0), cons1 = __any_sync(0xffffffffu, R.ncons[1] != 0.0);
            const bool cons2 = __any_sync(0xffffffffu, R.ncons[2] != 0.0);
            if (__any_sync(0xffffffffu, R.invc != 0.0)) R.variant = kCons0 | kCons2 | kInvc;
            else R.variant = (zero1 ? kZero1 : 0) | ((cons0 || (!zero1 && cons1)) ? kCons0 : 0) | (cons2 ? kCons2 : 0);
        }
        __syncwarp();
        pdl_wait();   // (the decision warps are gated by the aux warps' barriers anyway; this orders their V stores after the predecessors)
        if (q.trace && blockIdx.x == 0 && threadIdx.x == 0) q.trace[(0 * 64 + 63) * 8 + 7] = clock64();   // dependency wait over
        const Pipe4 P{sm, vfull, cready, dfree, rfull};
#if AB_NO_DECIDE
        decision_loop_empty(q, P, TP);
#else
        if ((LSMC_ABLATE == 13 && j >= 7) || (LSMC_ABLATE == 14 && (j & 3) != 0)) { decision_loop_empty(q, P, TP); return; }
        switch (R.variant) {
#define VARIANT(f) case f: decision_loop<TP, MTD, KS, f>(q, L, P, R, j, g, t); break;
            VARIANT(0) VARIANT(1) VARIANT(2) VARIANT(3) VARIANT(4) VARIANT(5) VARIANT(6) VARIANT(7)
            default: decision_loop<TP, MTD, KS, kCons0 | kCons2 | kInvc>(q, L, P, R, j, g, t); break;
#undef VARIANT
        }
#endif
    } else {
        // ============================================== aux warps ==============================================
        const int aw = warp - kDecWarps;
        const int alane = aw * 32 + lane;
        const int n_mt = GnP >> 3;                      // 8-row blocks of fitted values
        const uint32_t vn_bytes = (uint32_t)((size_t)q.Gn * TP * 8), raw_bytes = (uint32_t)(2 * (F + 1) * ROW);
        auto produce_vn = [&](int64_t tile, int b) {
            mbar_expect_tx(&vfull[b], vn_bytes);
            tma_2d(sm + L.vn + (size_t)b * L.vn_buf, &mapV, (int)(tile * TP), 0, &vfull[b]);
        };
        auto produce_raw = [&](int64_t tile, int s) {
            unsigned char* st = sm + L.raw + (size_t)s * L.stage;
            const int c0 = (int)(tile * TP);
            mbar_expect_tx(&rfull[s], raw_bytes);
            tma_2d(st + L.o_sp, &mapSp, c0, 0, &rfull[s]);
            tma_2d(st + L.o_fp, &mapFp, c0, 0, &rfull[s]);
            tma_2d(st + L.o_sm, &mapSm, c0, 0, &rfull[s]);
            tma_2d(st + L.o_fm, &mapFm, c0, 0, &rfull[s]);
        };
        // Product nodes of the basis monomials (both periods): when they form one wave of mutually independent
        // multiplies, every aux lane keeps its (period, node, path) elements as packed descriptors in registers
        // (operand offsets / 8 in 15 bits each + "relative to the raw stage" flags; destination byte offset), so a tile's
        // products are one round of independent loads, multiplies and stores.  Lanes without an element multiply the
        // zero row into itself.
        const int n_waves = sNWaves, tot_el = 2 * n_prod * TP;
        bool fastprod = n_waves <= 1 && tot_el <= 96 * kElemReg;
        for (int o = 0; o < n_prod; o++) fastprod = fastprod && !((sOpA[0][o] | sOpB[0][o]) & kOnes);
        uint32_t eAB[kElemReg], eD[kElemReg];
#pragma unroll
        for (int e = 0; e < kElemReg; e++) {
            const int idx = alane + 96 * e;
            uint32_t da = (uint32_t)L.zero, db = (uint32_t)L.zero, dd = (uint32_t)L.zero;
            if (fastprod && idx < tot_el) {
                const int pair = idx / TP, path = idx - pair * TP;
                const int sel = pair >= n_prod ? 1 : 0, o = pair - sel * n_prod;
                da = sOpA[sel][o] + 8u * path; db = sOpB[sel][o] + 8u * path; dd = sOpD[sel][o] + 8u * path;
            }
            eAB[e] = ((da & kOffMask) >> 3) | ((da & kRel) ? 0x8000u : 0u) | (((db & kOffMask) >> 3) << 16) | ((db & kRel) ? 0x80000000u : 0u);
            eD[e] = dd;
        }

        // One tile loop per role, so that the register-resident fragments of one role are not live in the other.
        auto run = [&](auto role) {
            constexpr int ROLE = decltype(role)::value;   // 0: warp 13 (copies, x~, Gram, last fit block); 1: warps 14, 15 (fit)
            // fit blocks: warp 14 -> 0..5, warp 15 -> 6..11, warp 13 -> 12
            constexpr int NB = ROLE == 0 ? 1 : kFitTiles;
            const int mt_base = ROLE == 0 ? 2 * kFitTiles : (aw - 1) * kFitTiles;
            const int mt_cnt = max(0, min(n_mt - mt_base, NB));
            // the first tile's regressor rows are inputs of the whole valuation (simulation panels): on their way before the
            // dependency wait.  Nothing else may be: under PDL "my predecessor has STARTED" is all a CTA knows until its own wait
            // returns, and when the grids are smaller than the GPU (few paths) the sweep before the tail kernel can still be
            // running — V_next, its output, must not be touched before the wait (a hoisted V_next copy read unfinished rows).
            if (ROLE == 0 && lane == 0 && (int64_t)blockIdx.x < n_tiles) produce_raw(blockIdx.x, 0);
            pdl_wait();   // the coefficients (tail kernel) and V_next (the sweep before it) are complete and visible from here on
            // beta fragments (DMMA A operand: A[g][t] = beta[8 mt + g][4 ks + t]) stay in registers for the whole launch
            double bfrag[NB][KSTEPS];
#pragma unroll
            for (int mi = 0; mi < NB; mi++)
#pragma unroll
                for (int ks = 0; ks < KSTEPS; ks++)
                    bfrag[mi][ks] = (mi < mt_cnt) ? __ldcg(q.beta + (size_t)(8 * (mt_base + mi) + g) * q.ldb + 4 * ks + t) : 0.0;
            // descriptors of the basis rows this lane feeds as DMMA B operand: B[t][g] = x[4 ks + t][8 nt + g]
            uint32_t descP[KSTEPS], descM[KSTEPS];
#pragma unroll
            for (int ks = 0; ks < KSTEPS; ks++) {
                const int k = 4 * ks + t;
                descP[ks] = (k < K) ? sNode[0][bd->node_of_basis[k]] : (uint32_t)L.zero;
                descM[ks] = (k < K && ROLE == 0) ? sNode[1][bd->node_of_basis[k]] : (uint32_t)L.zero;
            }
            // warp 13: T^T fragments (A[g][t] = T[4 ks + t][8 m + g]) and the Gram accumulators
            constexpr int MTX = ROLE == 0 ? MT : 1;
            double tfrag[MTX][KSTEPS];
            double xacc[MTX][MTX][2];
#pragma unroll
            for (int m = 0; m < MTX; m++) {
#pragma unroll
                for (int ks = 0; ks < KSTEPS; ks++) {
                    const int i2 = 4 * ks + t, j2 = 8 * m + g;
                    tfrag[m][ks] = (ROLE == 0 && i2 < K && j2 < K) ? __ldg(q.precond + (size_t)i2 * K + j2) : 0.0;
                }
#pragma unroll
                for (int n = 0; n < MTX; n++) xacc[m][n][0] = xacc[m][n][1] = 0.0;
            }

            int stage = 0, sphase = 0;
            int it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
                const int b = it & 1;
                long long* tr = (q.trace && blockIdx.x == 0 && it < 64 && lane == 0) ? q.trace + ((size_t)warp * 64 + it) * 8 : nullptr;
                if (tr) tr[0] = clock64();
                // buffers b (V_next, C^, x~) were last used by tile it-2: wait until every decision warp has left it
                if (it >= 2) mbar_wait(&dfree[b], (uint32_t)(((it >> 1) + 1) & 1));
                if (tr) tr[1] = clock64();
                if (ROLE == 0 && lane == 0) {
                    produce_vn(tile, b);
                    const int64_t nxt = tile + gridDim.x;
                    if (nxt < n_tiles) produce_raw(nxt, stage + 1 == kRawStages ? 0 : stage + 1);
                }
                mbar_wait(&rfull[stage], (uint32_t)sphase);
                if (tr) tr[2] = clock64();
                const uint32_t rawoff = (uint32_t)(L.raw + (size_t)stage * L.stage);
                const uint32_t onesoff = (tile == n_tiles - 1) ? (uint32_t)(L.ones_last - L.ones) : 0u;
                auto resolve = [&](uint32_t d) -> const double* {
                    return reinterpret_cast<const double*>(sm + (d & kOffMask) + ((d & kRel) ? rawoff : 0u) + ((d & kOnes) ? onesoff : 0u));
                };
                aux_bar();   // the previous tile's readers of the product rows are done
                if (tr) tr[5] = clock64();

                // ---- product nodes ----
                if (fastprod) {
                    double va[kElemReg], vb[kElemReg];
#pragma unroll
                    for (int e = 0; e < kElemReg; e++) {
                        const uint32_t w = eAB[e];
                        va[e] = *reinterpret_cast<const double*>(sm + ((w & 0x7fffu) << 3) + ((w & 0x8000u) ? rawoff : 0u));
                        vb[e] = *reinterpret_cast<const double*>(sm + (((w >> 16) & 0x7fffu) << 3) + ((w & 0x80000000u) ? rawoff : 0u));
                    }
#pragma unroll
                    for (int e = 0; e < kElemReg; e++) *reinterpret_cast<double*>(sm + eD[e]) = mul_rn(va[e], vb[e]);
                } else {
                    // general programs: wave by wave, the 96 aux lanes share the (period, node, path) elements of a wave
                    int o0 = 0;
                    for (int w = 0; w < n_waves; w++) {
                        // one (period, node) row per aux warp at a time: its descriptors are warp-uniform, lanes take the paths
                        const int o1 = sWaveEnd[w], nw = o1 - o0;
                        for (int pair = aw; pair < 2 * nw; pair += kAuxWarps) {
                            const int sel = pair >= nw ? 1 : 0, o = o0 + pair - sel * nw;
                            const double* pa = resolve(sOpA[sel][o]);
                            const double* pb = resolve(sOpB[sel][o]);
                            double* pd = reinterpret_cast<double*>(sm + sOpD[sel][o]);
#pragma unroll
                            for (int path = lane; path < TP; path += 32) pd[path] = mul_rn(pa[path], pb[path]);
                        }
                        o0 = o1;
                        if (w + 1 < n_waves) aux_bar();
                    }
                }
                if (tr) tr[6] = clock64();
                aux_bar();
                if (tr) tr[3] = clock64();

                double* sC = reinterpret_cast<double*>(sm + L.c) + (size_t)b * GnP * TP;
                const double* pp[KSTEPS];
#pragma unroll
                for (int ks = 0; ks < KSTEPS; ks++) pp[ks] = resolve(descP[ks]) + g;
                double* crow = sC + (size_t)(8 * mt_base + g) * TP + 2 * t;
                if constexpr (ROLE == 0) {
                    if (!AB_NO_XT) {
                        // ---- x~ = T^T x_{p-1} on the tensor pipe, then X~^T X~ ----
                        double* xt = reinterpret_cast<double*>(sm + L.xt) + (size_t)b * 8 * MT * TP;
                        const double* pm[KSTEPS];
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++) pm[ks] = resolve(descM[ks]) + g;
                        double bnx[KSTEPS];   // B operands one step ahead of their use
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++) bnx[ks] = pm[ks][0];
#pragma unroll
                        for (int nt = 0; nt < TP / 8; nt++) {
                            double bop[KSTEPS], cx[MT][2];
#pragma unroll
                            for (int ks = 0; ks < KSTEPS; ks++) { bop[ks] = bnx[ks]; if (nt + 1 < TP / 8) bnx[ks] = pm[ks][8 * (nt + 1)]; }
#pragma unroll
                            for (int m = 0; m < MT; m++) cx[m][0] = cx[m][1] = 0.0;
#pragma unroll
                            for (int ks = 0; ks < KSTEPS; ks++)
#pragma unroll
                                for (int m = 0; m < MT; m++) dmma8x8x4(cx[m][0], cx[m][1], tfrag[m][ks], bop[ks]);
#pragma unroll
                            for (int m = 0; m < MT; m++)
                                *reinterpret_cast<double2*>(xt + (size_t)(8 * m + g) * TP + 8 * nt + 2 * t) = make_double2(cx[m][0], cx[m][1]);
                        }
                        __syncwarp();
                        const double* px = xt + (size_t)g * TP + 2 * t;
#pragma unroll
                        for (int u8 = 0; u8 < TP; u8 += 8) {
                            double2 f2[MT];
#pragma unroll
                            for (int m = 0; m < MT; m++) f2[m] = lds128(px + (size_t)8 * m * TP + u8);
#pragma unroll
                            for (int m = 0; m < MT; m++)
#pragma unroll
                                for (int n = 0; n < MT; n++) {
                                    dmma8x8x4(xacc[m][n][0], xacc[m][n][1], f2[m].x, f2[n].x);
                                    dmma8x8x4(xacc[m][n][0], xacc[m][n][1], f2[m].y, f2[n].y);
                                }
                        }
                    }
                    // ---- the last block of fitted values ----
                    if (!AB_NO_FIT && mt_cnt > 0) {
                        double bnx[KSTEPS];
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++) bnx[ks] = pp[ks][0];
#pragma unroll
                        for (int nt = 0; nt < TP / 8; nt++) {
                            double bop[KSTEPS], c0 = 0.0, c1 = 0.0;
#pragma unroll
                            for (int ks = 0; ks < KSTEPS; ks++) { bop[ks] = bnx[ks]; if (nt + 1 < TP / 8) bnx[ks] = pp[ks][8 * (nt + 1)]; }
#pragma unroll
                            for (int ks = 0; ks < KSTEPS; ks++) dmma8x8x4(c0, c1, bfrag[0][ks], bop[ks]);
                            *reinterpret_cast<double2*>(crow + 8 * nt) = make_double2(c0, c1);
                        }
                    }
                } else if (!AB_NO_FIT) {
                    // ---- fitted continuation values C^[l][n] = beta_l . x_p[n] on the FP64 tensor pipe: six row blocks per
                    //      warp, independent accumulator chains issued back to back, stores predicated ----
                    double bnx[KSTEPS];
#pragma unroll
                    for (int ks = 0; ks < KSTEPS; ks++) bnx[ks] = pp[ks][0];
#pragma unroll
                    for (int nt = 0; nt < TP / 8; nt++) {
                        double bop[KSTEPS], cf[NB][2];
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++) { bop[ks] = bnx[ks]; if (nt + 1 < TP / 8) bnx[ks] = pp[ks][8 * (nt + 1)]; }
#pragma unroll
                        for (int mi = 0; mi < NB; mi++) cf[mi][0] = cf[mi][1] = 0.0;
#pragma unroll
                        for (int ks = 0; ks < KSTEPS; ks++)
#pragma unroll
                            for (int mi = 0; mi < NB; mi++) dmma8x8x4(cf[mi][0], cf[mi][1], bfrag[mi][ks], bop[ks]);
#pragma unroll
                        for (int mi = 0; mi < NB; mi++)
                            if (mi < mt_cnt) *reinterpret_cast<double2*>(crow + (size_t)8 * mi * TP + 8 * nt) = make_double2(cf[mi][0], cf[mi][1]);
                    }
                }
                if (tr) tr[4] = clock64();
                mbar_arrive(&cready[b]);
                if (++stage == kRawStages) { stage = 0; sphase ^= 1; }
            }
            if constexpr (ROLE == 0) {
                double* po = q.partial + (size_t)blockIdx.x * (8 * MT) * q.ncols;
#pragma unroll
                for (int m = 0; m < MT; m++)
#pragma unroll
                    for (int n = 0; n < MT; n++)
                        *reinterpret_cast<double2*>(po + (size_t)(8 * m + g) * q.ncols + 8 * (NT + n) + 2 * t) = make_double2(xacc[m][n][0], xacc[m][n][1]);
            }
        };
        if (aw == 0) run(std::integral_constant<int, 0>{});
        else run(std::integral_constant<int, 1>{});
    }
}

// ------------------------------------------------------------------------------------------------
// Launch
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// [rows][Npad] FP64 panel, box = TP paths x box_rows rows.
bool make_panel_map(CUtensorMap_st* map, const double* base, int64_t Npad, int rows, int TP, int box_rows) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)Npad, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)Npad * 8};
    const cuuint32_t box[2] = {(cuuint32_t)TP, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Paths per tile for the pipelined kernel, 0 when the problem's shape is outside what it covers.
int sweep4_tile(int max_grid, int K, int F, int n_prod) {
    if (K > 16 || max_grid > 8 * kDecWarps || F > kMaxFactors || n_prod > kMaxProd) return 0;
    if (!encode_tiled_fn()) return 0;
    const size_t limit = 227 * 1024 - 2048;   // dynamic shared memory + the kernel's static part (barriers, descriptors)
    const int tps[2] = {56, 40};
    for (int tp : tps)
        if (layout4(tp, max_grid, K, F, n_prod).total <= limit) return tp;
    return 0;
}

template <int TP, int MTD, int KS, int KSTEPS>
static cudaError_t launch4_t(SweepParams& q, int n_prod, int blocks, cudaStream_t stream) {
    auto kern = sweep4_kernel<TP, MTD, KS, KSTEPS>;
    const size_t smem = layout4(TP, q.Gn, q.K, q.F, n_prod).total;
    q.n_prod = n_prod;
    if (cudaError_t e = ensure_dynamic_smem((const void*)kern, smem)) return e;
    CUtensorMap mv, msp, mfp, msm, mfm;
    if (!make_panel_map(&mv, q.Vnext, q.Npad, q.Gn, TP, q.Gn) || !make_panel_map(&msp, q.spot_p, q.Npad, 1, TP, 1) ||
        !make_panel_map(&mfp, q.fac_p, q.Npad, q.F, TP, q.F) || !make_panel_map(&msm, q.spot_m, q.Npad, 1, TP, 1) ||
        !make_panel_map(&mfm, q.fac_m, q.Npad, q.F, TP, q.F))
        return cudaErrorInvalidValue;
    return launch_pdl(q.pdl != 0, kern, blocks, kWarps * 32, smem, stream, q, mv, msp, mfp, msm, mfm);
}

template <int TP>
static cudaError_t launch4_tp(SweepParams& q, int n_prod, int blocks, cudaStream_t stream) {
    const int K = q.K;
    // whole blocks of 8 basis rows go through DMMA, a remainder of 1..3 rows through DFMA
#ifdef LSMC_ALL_DMMA   // development: remainder rows through a padded DMMA block as well
    const int MTD = (K + 7) / 8;
#else
    const int MTD = (K % 8 != 0 && K % 8 <= 3) ? K / 8 : (K + 7) / 8;
#endif
    const int KS = K - 8 * MTD > 0 ? K - 8 * MTD : 0;
    const int KSTEPS = (K + 3) / 4;
#define SHAPE(m, s, k) if (MTD == m && KS == s && KSTEPS == k) return launch4_t<TP, m, s, k>(q, n_prod, blocks, stream);
#ifdef LSMC_FAST_BUILD
    SHAPE(0, 3, 1) SHAPE(1, 2, 3) SHAPE(1, 0, 1) SHAPE(2, 0, 3)
#else
    SHAPE(0, 1, 1) SHAPE(0, 2, 1) SHAPE(0, 3, 1) SHAPE(1, 0, 1) SHAPE(1, 0, 2) SHAPE(1, 1, 3) SHAPE(1, 2, 3) SHAPE(1, 3, 3)
    SHAPE(2, 0, 3) SHAPE(2, 0, 4)
#endif
#undef SHAPE
    return cudaErrorInvalidValue;
}

// Regular period only: decide mode, fitted continuation values, normal equations of period p-1, simulated spot.
bool sweep4_covers(const SweepParams& q) {
    return q.mode == 0 && q.cont_mode == 0 && q.accum_mode == 1 && !q.spot_is_const && q.K <= 16 && q.G <= 8 * kDecWarps &&
           q.Gn <= 8 * kDecWarps && q.max_dec <= 3;
}

cudaError_t launch_sweep4(SweepParams& q, int TP, int n_prod, int blocks, cudaStream_t stream) {
    q.ncols = 8 * ((q.G + 7) / 8 + (q.K + 7) / 8);
    if (TP == 56) return launch4_tp<56>(q, n_prod, blocks, stream);
#ifndef LSMC_FAST_BUILD
    if (TP == 40) return launch4_tp<40>(q, n_prod, blocks, stream);
#endif
    return cudaErrorInvalidValue;
}

}  // namespace lsmc
