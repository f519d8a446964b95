// sweep.cu — K3: the fused backward sweep of one period (LSV.cs:185-205, 234-332, 753-770), sm_100a.
//
// One persistent CTA per SM walks path tiles of TP paths.  16 warps:
//   warps 0..14  "workers": fit (DMMA), decisions (8 grid rows x TP paths per item) and X~^T X items
//   warp  15     "producer" during the decision phase: bulk-async copies (cp.async.bulk -> UBLKCP, mbarrier
//                completion) of the NEXT tile's V_next rows and raw regressor rows (spot, factors of periods p and
//                p-1) into the other shared-memory buffer, so HBM latency hides behind the FP64 work of this tile.
// Per tile:
//   F0 basis monomials of both periods from the raw rows (all threads, one (basis, path) element each)
//   F1 x~ = T^T x_{p-1} (preconditioned regressors) and the fitted continuation values C^[l][n] = beta_l . x_p[n]
//      for every next-grid row l on the FP64 tensor pipe (DMMA 8x8x4: A = beta, B = X_p^T) -> shared memory
//   D  every warp first takes its share of the 8 x 8 blocks of X~^T X~ (lower block triangle, dealt round robin); then
//      lane (g,t) of a decision item owns grid row i = 8j+g and, per iteration, the two paths 8u+2t, 8u+2t+1
//      (128-bit shared loads): max over the bang-bang decisions of immediate + interpolated C^, realised value from
//      interpolated V_next, store V[i][n] (128-bit), and feed the value straight from registers as the B operand
//      of the DMMA that accumulates X~_{p-1}^T V.
// All arithmetic that decides is individually rounded (__dmul_rn/__dadd_rn) in the reference's order: given the same
// coefficients the value surface is bit-identical to the oracle's.  FP64 scalar and DMMA share one pipe on B200
// (scripts/probe_fp64.cu: 36 TFLOP/s either way, not additive), which — not HBM — is what bounds this kernel.
#include <cuda.h>

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
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    const uint32_t addr = smem_u32(bar);
    while (!done) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    }
}
// 1-D bulk asynchronous copy global -> shared (TMA unit, no tensor map), completion on an mbarrier.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// 2-D tensor copy global -> shared through a TMA descriptor: the whole V_next tile in one instruction (a bulk copy per row —
// a hundred 192-byte copies per tile — kept the producer warp issuing for most of a tile's time).
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}

// Shared-space accesses by 32-bit address (the basis program indexes the whole carve-up with offsets from its op table)
__device__ __forceinline__ double lds64_at(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts64_at(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ int4 lds128i_at(uint32_t a) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}

__device__ __forceinline__ double2 lds128(const double* p) { return *reinterpret_cast<const double2*>(p); }

#ifndef LSMC_DECIDE_UNROLL
#define LSMC_DECIDE_UNROLL 3
#endif
constexpr int kDecideUnroll = LSMC_DECIDE_UNROLL;
constexpr int kSweepWarps = 16;
constexpr int kWorkers = 15;

// Shared-memory carve-up (in doubles).  Row strides: LD = TP (TP = 8 mod 16 keeps the 128-bit row-pair accesses of
// the decision loop conflict-free), LDX = TP + 4 for the design rows read as DMMA B operands, LDB for beta rows.
struct SweepSmem {
    int LD, LDX, LDB, GnP, Kp4, nbuf, raw_rows;
    size_t vn, raw, c, xp, xm, xt, beta, t, nodes, cst, ops, total;   // offsets in doubles
};

__host__ __device__ inline SweepSmem sweep_layout(int TP, int Gn, int K, int F, int MT, int nbuf, int n_scratch, int n_rowops) {
    SweepSmem L;
    L.LD = TP; L.LDX = TP + 4; L.nbuf = nbuf;
    L.GnP = (Gn + 7) & ~7;
    L.Kp4 = (K + 3) & ~3;
    L.LDB = L.Kp4 + ((L.Kp4 % 8 == 0) ? 4 : 0);
    L.raw_rows = 2 * (F + 1);
    size_t o = 0;
    L.vn = o;    o += (size_t)nbuf * L.GnP * L.LD;           // staged V_next tile(s), filled by bulk copies
    L.raw = o;   o += (size_t)nbuf * L.raw_rows * L.LD;      // raw regressors: spot_p, fac_p[F], spot_{p-1}, fac_{p-1}[F]
    L.c = o;     o += (size_t)L.GnP * L.LD;                  // fitted continuation values
    L.xp = o;    o += (size_t)L.Kp4 * L.LDX;                 // basis rows of period p (rows >= K stay zero)
    L.xm = o;    o += (size_t)L.Kp4 * L.LDX;                 // basis rows of period p-1 (rows >= K stay zero)
    L.xt = o;    o += (size_t)nbuf * 8 * MT * L.LD;          // preconditioned rows x~ (rows >= K stay zero), one set per tile buffer
    L.beta = o;  o += (size_t)L.GnP * L.LDB;
    L.t = o;     o += (size_t)L.Kp4 * (8 * MT + 4);          // preconditioner, row stride 8 MT + 4 (rows / columns >= K stay zero)
    L.nodes = o; o += (size_t)2 * n_scratch * L.LD;              // product nodes that are not basis functions, both periods
    L.cst = o;   o += (size_t)2 * L.LD;                          // row 0 zeros, row 1 ones (zero beyond the last path)
    o = (o + 1) & ~(size_t)1;
    L.ops = o;   o += (size_t)2 * n_rowops * 2;                  // the row program of both periods as int4 {A, B, D offsets, raw flags}
    L.total = o;
    return L;
}

// Decision table of one grid row, register resident for the whole kernel (the first three decisions of the
// bang-bang set; rows with fewer repeat decision 0, which a strict '>' never selects; more go through global memory).
struct RowTable {
    int nd, row_ok, variant;
    double invc, gridv;
    double nvol[3], cost[3], ncons[3], wl[3], wh[3];   // nvol = -volume, ncons = -consumed (exact negations)
    int oL[3], oH[3];
};

// Loop variants (warp-uniform, chosen once per item): which exactly-neutral operations can be skipped.
enum : int { kZero1 = 1, kCons0 = 2, kCons2 = 4, kInvc = 8, kGeneric = 16 };

// One decision item of one tile: 8 grid rows (lane group g) x TP paths; lane (g,t) handles paths 8u+2t, 8u+2t+1.
//   tot_d = ((-vol_d s) df - cost_d [+ (-cons_d s) df]) + cont_d [- invc]      (LSV.cs:304-330, operation order kept)
// Skipped terms are exact no-ops: adding (-0 * s) * df = -0.0, subtracting invc = +0.0, and for the zero decision
// (vol = cost = cons = 0) the immediate value is -0.0, so tot = cont bit for bit.
template <int TP, int MT, int FLAGS>
__device__ __forceinline__ void decide_item(const RowTable& R, const double* __restrict__ sC, const double* __restrict__ vn,
                                            const double* __restrict__ sS, const double* __restrict__ sXT, double* __restrict__ vrow,
                                            const double df, const bool accum, const int g, const int t, double (&acc)[MT][2]) {
    constexpr bool ZERO1 = (FLAGS & kZero1) != 0, CONS0 = (FLAGS & kCons0) != 0, CONS2 = (FLAGS & kCons2) != 0, INVC = (FLAGS & kInvc) != 0;
    constexpr bool CONS1 = CONS0 || CONS2;
    constexpr int LD = TP;
    const double* pS = sS + 2 * t;
    const double* pL0 = sC + R.oL[0] + 2 * t; const double* pH0 = sC + R.oH[0] + 2 * t;
    const double* pL1 = sC + R.oL[1] + 2 * t; const double* pH1 = sC + R.oH[1] + 2 * t;
    const double* pL2 = sC + R.oL[2] + 2 * t; const double* pH2 = sC + R.oH[2] + 2 * t;
    const double* pV = vn + 2 * t;
    const double* pX = sXT + (size_t)g * LD + 2 * t;
    double* out = vrow + 2 * t;
#pragma unroll (kDecideUnroll)   // 3: the whole 24-path tile of the K > 16 shapes straight-line (no spills there; 1.357 -> 1.320 ms at C4)
    for (int u = 0; u < TP / 8; u++) {
        double v[2] = {0.0, 0.0};
        if (R.row_ok) {
            const double2 s2 = lds128(pS + 8 * u);
            const double2 l0 = lds128(pL0 + 8 * u), h0 = lds128(pH0 + 8 * u);
            const double2 l1 = lds128(pL1 + 8 * u), h1 = lds128(pH1 + 8 * u);
            const double2 l2 = lds128(pL2 + 8 * u), h2 = lds128(pH2 + 8 * u);
            const double sv[2] = {s2.x, s2.y};
            const double cl[3][2] = {{l0.x, l0.y}, {l1.x, l1.y}, {l2.x, l2.y}};
            const double ch[3][2] = {{h0.x, h0.y}, {h1.x, h1.y}, {h2.x, h2.y}};
#pragma unroll
            for (int e = 0; e < 2; e++) {
                double cont[3], tot[3];
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    cont[d] = add_rn(mul_rn(ch[d][e], R.wh[d]), mul_rn(cl[d][e], R.wl[d]));
                    double x;
                    if (d == 1 && ZERO1) x = cont[d];
                    else {
                        double imm = sub_rn(mul_rn(mul_rn(R.nvol[d], sv[e]), df), R.cost[d]);
                        if ((d == 0 && CONS0) || (d == 1 && CONS1) || (d == 2 && CONS2)) imm = add_rn(imm, mul_rn(mul_rn(R.ncons[d], sv[e]), df));
                        x = add_rn(imm, cont[d]);
                    }
                    tot[d] = INVC ? sub_rn(x, R.invc) : x;
                }
                double best = tot[0], bestC = cont[0];
                int kb = 0;
                if (tot[1] > best) { best = tot[1]; bestC = cont[1]; kb = 1; }
                if (tot[2] > best) { best = tot[2]; bestC = cont[2]; kb = 2; }
                const int oL = kb == 0 ? R.oL[0] : (kb == 1 ? R.oL[1] : R.oL[2]);
                const int oH = kb == 0 ? R.oH[0] : (kb == 1 ? R.oH[1] : R.oH[2]);
                const double wl = kb == 0 ? R.wl[0] : (kb == 1 ? R.wl[1] : R.wl[2]);
                const double wh = kb == 0 ? R.wh[0] : (kb == 1 ? R.wh[1] : R.wh[2]);
                const double act = add_rn(mul_rn(pV[oH + 8 * u + e], wh), mul_rn(pV[oL + 8 * u + e], wl));
                v[e] = add_rn(best, add_rn(-bestC, act));
            }
            __stcs(reinterpret_cast<double2*>(out + 8 * u), make_double2(v[0], v[1]));
        }
        if (accum) {
#pragma unroll
            for (int m = 0; m < MT; m++) {
                const double2 a2 = lds128(pX + (size_t)8 * m * LD + 8 * u);
                dmma8x8x4(acc[m][0], acc[m][1], a2.x, v[0]);
                dmma8x8x4(acc[m][0], acc[m][1], a2.y, v[1]);
            }
        }
    }
}

// Generic item: any number of decisions per row (extra_decisions > 0), decision table read through L1.
template <int TP, int MT>
__device__ __noinline__ void decide_item_generic(const SweepParams& q, const RowTable& R, const int i, const double* __restrict__ sC,
                                                 const double* __restrict__ vn, const double* __restrict__ sS, const double* __restrict__ sXT,
                                                 double* __restrict__ vrow, const bool accum, const int g, const int t, double (&acc)[MT][2]) {
    constexpr int LD = TP;
    const size_t trow = (size_t)(R.row_ok ? i : 0) * q.max_dec;
    for (int u = 0; u < TP / 8; u++) {
        const int pc = 8 * u + 2 * t;
        double v[2] = {0.0, 0.0};
        if (R.row_ok) {
            for (int e = 0; e < 2; e++) {
                const double s = sS[pc + e];
                double best = 0.0, bestC = 0.0, bwl = 1.0, bwh = 0.0;
                int bl = 0, bh = 0;
                for (int d = 0; d < R.nd; d++) {
                    const int l = __ldg(q.t_lo + trow + d) * LD, h = __ldg(q.t_hi + trow + d) * LD;
                    const double wlo = __ldg(q.t_wlo + trow + d), whi = __ldg(q.t_whi + trow + d);
                    const double c = add_rn(mul_rn(sC[h + pc + e], whi), mul_rn(sC[l + pc + e], wlo));
                    const double a = mul_rn(mul_rn(-__ldg(q.t_vol + trow + d), s), q.df);
                    const double cc = mul_rn(mul_rn(-__ldg(q.t_cons + trow + d), s), q.df);
                    const double x = sub_rn(add_rn(add_rn(sub_rn(a, __ldg(q.t_cost + trow + d)), cc), c), R.invc);
                    if (d == 0 || x > best) { best = x; bestC = c; bl = l; bh = h; bwl = wlo; bwh = whi; }
                }
                const double act = add_rn(mul_rn(vn[bh + pc + e], bwh), mul_rn(vn[bl + pc + e], bwl));
                v[e] = add_rn(best, add_rn(-bestC, act));
            }
            __stcs(reinterpret_cast<double2*>(vrow + pc), make_double2(v[0], v[1]));
        }
        if (accum) {
#pragma unroll
            for (int m = 0; m < MT; m++) {
                const double2 a2 = lds128(sXT + (size_t)(8 * m + g) * LD + pc);
                dmma8x8x4(acc[m][0], acc[m][1], a2.x, v[0]);
                dmma8x8x4(acc[m][0], acc[m][1], a2.y, v[1]);
            }
        }
    }
}

// Terminal period (LSV.cs:118-128): V[i][n] = TerminalStorageNpv(spot_end[n], grid_end[i]).
template <int TP, int MT>
__device__ __forceinline__ void terminal_item(const SweepParams& q, const RowTable& R, const double* __restrict__ sS,
                                              const double* __restrict__ sXT, double* __restrict__ vrow, const bool accum,
                                              const int g, const int t, double (&acc)[MT][2]) {
    constexpr int LD = TP;
#pragma unroll 1
    for (int u = 0; u < TP / 8; u++) {
        const int pc = 8 * u + 2 * t;
        double v[2] = {0.0, 0.0};
        if (R.row_ok) {
            const double2 s2 = lds128(sS + pc);
            const double sv[2] = {s2.x, s2.y};
#pragma unroll
            for (int e = 0; e < 2; e++)
                v[e] = q.must_be_empty ? 0.0
                    : add_rn(add_rn(add_rn(q.tc0, mul_rn(q.tc1, sv[e])), mul_rn(q.tc2, R.gridv)), mul_rn(q.tc3, mul_rn(sv[e], R.gridv)));
            __stcs(reinterpret_cast<double2*>(vrow + pc), make_double2(v[0], v[1]));
        }
        if (accum) {
#pragma unroll
            for (int m = 0; m < MT; m++) {
                const double2 a2 = lds128(sXT + (size_t)(8 * m + g) * LD + pc);
                dmma8x8x4(acc[m][0], acc[m][1], a2.x, v[0]);
                dmma8x8x4(acc[m][0], acc[m][1], a2.y, v[1]);
            }
        }
    }
}

}  // namespace

template <int TP, int MT, int NTW>
__global__ void __launch_bounds__(kSweepWarps * 32, 1) sweep_kernel(SweepParams q, int nbuf, const __grid_constant__ CUtensorMap mapV) {
    extern __shared__ __align__(128) double smem[];
    pdl_launch_dependents();   // PDL: see lsmc_device.cuh
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const BasisDev* __restrict__ bd = q.basis;
    const int K = q.K, F = q.F;
    const int n_rowops = bd->n_rowops, n_scr = bd->n_scratch;
    const SweepSmem L = sweep_layout(TP, q.Gn, K, F, MT, nbuf, n_scr, n_rowops);
    constexpr int LD = TP, LDX = TP + 4;
    constexpr int LDT = 8 * MT + 4;   // strides = 4 mod 8 doubles: the four k-rows of a DMMA operand fall into distinct bank groups
    const int LDB = L.LDB, GnP = L.GnP, Kp4 = L.Kp4;
    double* const sVn = smem + L.vn;
    double* const sRaw = smem + L.raw;
    double* const sC = smem + L.c;
    double* const sXP = smem + L.xp;
    double* const sXM = smem + L.xm;
    double* const sXT0 = smem + L.xt;
    double* const sBeta = smem + L.beta;
    double* const sT = smem + L.t;
    int4* const sOp4 = reinterpret_cast<int4*>(smem + L.ops);
    __shared__ uint64_t s_full[2];
    __shared__ int16_t sWaveEnd[kMaxRowOps];

    const int NT = (q.G + 7) >> 3;                        // decision items (8 grid rows each)
    const int n_items = NT;
    const int64_t n_tiles = (q.N + TP - 1) / TP;
    const bool need_sp = !q.spot_is_const;                          // spot of period p (else the deterministic price)
    const bool need_fp = q.mode == 0 && q.cont_mode == 0;           // basis of period p for the fitted values
    const bool need_m = q.accum_mode == 1;                          // basis of period p-1 for the normal equations
    const bool accum = q.accum_mode != 0;

    // ---- one-time staging ----
    const int n_waves = bd->n_rowwaves;
    for (int w = tid; w < n_waves; w += blockDim.x) sWaveEnd[w] = bd->rowwave_end[w];
    for (size_t i = tid; i < L.total; i += blockDim.x) smem[i] = 0.0;
    __syncthreads();
    // The row program (host_plan.cpp compile_row_program) of both periods with shared-memory offsets: per wave the entries of
    // period p, then those of period p-1.  Operands in the raw rows move with the tile buffer (flags: bit 0 A, bit 1 B).
    for (int e = tid; e < 2 * n_rowops; e += blockDim.x) {
        const int sel = e >= n_rowops ? 1 : 0, o = e - sel * n_rowops;
        int o0 = 0, o1 = 0;
        for (int w = 0; w < n_waves; w++) { o1 = bd->rowwave_end[w]; if (o < o1) break; o0 = o1; }
        const RowOpDev op = bd->rowops[o];
        auto offset = [&](int region, int row) -> int {
            if (region == 0) return (int)(L.cst + (size_t)row * LD);
            if (region == 1) return (int)(L.raw + (size_t)(sel * (F + 1) + row) * LD);
            if (region == 2) return (int)((sel ? L.xm : L.xp) + (size_t)row * LDX);
            return (int)(L.nodes + (size_t)(sel * n_scr + row) * LD);
        };
        sOp4[2 * o0 + sel * (o1 - o0) + (o - o0)] =
            make_int4(offset(op.ra, op.a), offset(op.rb, op.b), offset(op.rd, op.d), (op.ra == 1 ? 1 : 0) | (op.rb == 1 ? 2 : 0));
    }
    for (int path = tid; path < TP; path += blockDim.x) {
        smem[L.cst + LD + path] = 1.0;
        if (!need_sp) { sRaw[path] = q.spot_const; if (nbuf == 2) sRaw[(size_t)L.raw_rows * LD + path] = q.spot_const; }   // every tile buffer
    }
    pdl_wait();   // everything above reads only per-problem constants; the coefficients and V_next come from the previous kernels
    if (need_fp) {
        for (int i = tid; i < q.Gn * Kp4; i += blockDim.x) {
            const int l = i / Kp4, k = i - l * Kp4;
            sBeta[(size_t)l * LDB + k] = __ldg(q.beta + (size_t)l * q.ldb + k);
        }
    }
    if (need_m) for (int i = tid; i < K * K; i += blockDim.x) sT[(i / K) * LDT + i % K] = __ldg(q.precond + i);
    if (tid == 0) {
        mbar_init(&s_full[0], 1);
        mbar_init(&s_full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }

    // ---- decision tables -> registers ----
    RowTable tab[NTW];
    double acc[NTW][MT][2];
    // X~^T X~: the MT (MT + 1) / 2 blocks (kb, m), m <= kb, of the lower block triangle — block (m, kb) is its transpose —
    // dealt over the fifteen workers (the producer warp needs most of the phase to issue its copies); block
    // b = kb (kb + 1) / 2 + m goes to worker b % 15, slot b / 15
    constexpr int kGramBlocks = MT * (MT + 1) / 2, GB = (kGramBlocks + kWorkers - 1) / kWorkers;
    double gacc[GB][2];
    int gblk[GB];   // kb << 8 | m, -1: none
#pragma unroll
    for (int sl = 0; sl < GB; sl++) {
        gacc[sl][0] = gacc[sl][1] = 0.0;
        const int bidx = warp + sl * kWorkers;
        int kb = 0;
        while ((kb + 1) * (kb + 2) / 2 <= bidx) kb++;
        gblk[sl] = (q.accum_mode && warp < kWorkers && bidx < kGramBlocks) ? ((kb << 8) | (bidx - kb * (kb + 1) / 2)) : -1;
    }
#pragma unroll
    for (int jj = 0; jj < NTW; jj++) {
#pragma unroll
        for (int m = 0; m < MT; m++) acc[jj][m][0] = acc[jj][m][1] = 0.0;
        RowTable& R = tab[jj];
        const int j = warp + jj * kWorkers;
        const int i = 8 * j + g;
        R.row_ok = (warp < kWorkers && j < NT && i < q.G) ? 1 : 0;
        R.nd = 0; R.invc = 0.0; R.gridv = 0.0;
        const bool on = q.mode == 0 && R.row_ok;
        const size_t trow = (size_t)(R.row_ok ? i : 0) * q.max_dec;
        if (on) { R.nd = __ldg(q.t_ndec + i); R.invc = __ldg(q.t_invcost + i); }
        if (q.mode == 1 && R.row_ok) R.gridv = __ldg(q.grid + i);
#pragma unroll
        for (int d = 0; d < 3; d++) {
            const int ds = (d < R.nd) ? d : 0;   // missing decisions repeat decision 0
            R.nvol[d] = on ? -__ldg(q.t_vol + trow + ds) : 0.0;
            R.cost[d] = on ? __ldg(q.t_cost + trow + ds) : 0.0;
            R.ncons[d] = on ? -__ldg(q.t_cons + trow + ds) : 0.0;
            R.wl[d] = on ? __ldg(q.t_wlo + trow + ds) : 1.0;
            R.wh[d] = on ? __ldg(q.t_whi + trow + ds) : 0.0;
            R.oL[d] = on ? __ldg(q.t_lo + trow + ds) * LD : 0;
            R.oH[d] = on ? __ldg(q.t_hi + trow + ds) * LD : 0;
        }
        const bool zero1 = __all_sync(0xffffffffu, R.nvol[1] == 0.0 && R.cost[1] == 0.0 && R.ncons[1] == 0.0);
        const bool cons0 = __any_sync(0xffffffffu, R.ncons[0] != 0.0), cons1 = __any_sync(0xffffffffu, R.ncons[1] != 0.0);
        const bool cons2 = __any_sync(0xffffffffu, R.ncons[2] != 0.0);
        if (__any_sync(0xffffffffu, R.nd > 3)) R.variant = kGeneric;
        else if (__any_sync(0xffffffffu, R.invc != 0.0)) R.variant = kCons0 | kCons2 | kInvc;
        else R.variant = (zero1 ? kZero1 : 0) | ((cons0 || (!zero1 && cons1)) ? kCons0 : 0) | (cons2 ? kCons2 : 0);
    }
    __syncthreads();

    // ---- producer: bulk-async copies of one tile's V_next rows and raw regressor rows into buffer b ----
    auto produce = [&](int64_t tile, int b) {
        const int64_t n0 = tile * TP;
        const uint32_t row_bytes = (uint32_t)(TP * sizeof(double));
        const int nrows = (q.mode == 0 ? q.Gn : 0) + (need_sp ? 1 : 0) + (need_fp ? F : 0) + (need_m ? F + 1 : 0);
        if (nrows == 0) return;
        if (lane == 0) mbar_expect_tx(&s_full[b], (uint32_t)nrows * row_bytes);
        __syncwarp();
        if (q.mode == 0 && lane == 0) tma_2d(sVn + (size_t)b * GnP * LD, &mapV, (int)n0, 0, &s_full[b]);   // Gn rows x TP paths, row pitch LD = TP
        if (lane < 2 * (F + 1)) {
            const int r = lane;
            const double* src = nullptr;
            if (r == 0) { if (need_sp) src = q.spot_p + n0; }
            else if (r <= F) { if (need_fp) src = q.fac_p + (size_t)(r - 1) * q.Npad + n0; }
            else if (r == F + 1) { if (need_m) src = q.spot_m + n0; }
            else if (need_m) src = q.fac_m + (size_t)(r - F - 2) * q.Npad + n0;
            if (src) bulk_g2s(sRaw + ((size_t)b * L.raw_rows + r) * LD, src, row_bytes, &s_full[b]);
        }
    };
    const bool any_copy = q.mode == 0 || need_sp || need_m;

    // ---- phase F0 as a routine of `nth` cooperating threads (this thread is number `tix`; `sync` is their barrier): masks the
    //      columns beyond the last path, then runs the basis program on tile buffer bb.
    //      The basis program (host_plan.cpp compile_basis: left-to-right product chains, spot first, factors ascending, shared
    //      prefixes computed once — the order BasisFunctions.cs:92-132 and the oracle multiply in), one multiply per element,
    //      wave by wave; a node that is a basis function lands in its design-matrix row. ----
    auto basis_rows = [&](const int bb, const int64_t tn0, const int tix, const int nth, auto sync) {
        const int rb = bb * L.raw_rows * LD;
        if (tn0 + TP > q.N) {   // the last tile: columns beyond the last path must contribute nothing
            for (int path = tix; path < TP; path += nth) if (tn0 + path >= q.N) smem[L.cst + LD + path] = 0.0;
            for (int e = tix; e < L.raw_rows * TP; e += nth) {
                const int r = e / TP, path = e - r * TP;
                if (tn0 + path >= q.N) sRaw[rb + (size_t)r * LD + path] = 0.0;
            }
            sync();
        }
        int o0 = 0;
        for (int w = 0; w < n_waves; w++) {
            const int o1 = sWaveEnd[w], nw = o1 - o0;
            const int lo = need_fp ? 0 : nw, hi = need_m ? 2 * nw : nw;
            const uint32_t ops_a = smem_u32(sOp4 + 2 * o0 + lo), sm_a = smem_u32(smem);
            // four elements per step, all their loads ahead of the stores (the elements of a wave are independent, which the
            // compiler cannot know: one element per step ran at one shared-memory round trip after the other)
            constexpr int U = 4;
            const int n_el = (hi - lo) * TP;
            for (int e0 = tix; e0 < n_el; e0 += U * nth) {
                double va[U], vb[U];
                int4 op[U];
                int path[U];
                // (an element beyond the end repeats the wave's first one: same value stored twice, no branches)
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int e = e0 + u * nth < n_el ? e0 + u * nth : 0;
                    const int idx = e / TP;
                    path[u] = e - idx * TP;
                    op[u] = lds128i_at(ops_a + 16 * idx);
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    va[u] = lds64_at(sm_a + 8 * (op[u].x + ((op[u].w & 1) ? rb : 0) + path[u]));
                    vb[u] = lds64_at(sm_a + 8 * (op[u].y + ((op[u].w & 2) ? rb : 0) + path[u]));
                }
#pragma unroll
                for (int u = 0; u < U; u++) sts64_at(sm_a + 8 * (op[u].z + path[u]), mul_rn(va[u], vb[u]));
            }
            o0 = o1;
            sync();
        }
    };
    constexpr int NNT = TP / 8;
    // x~ = T^T x_{p-1}: D[j][path] = sum_i T[i][j] x[i][path]; the k-steps alternate between two accumulator pairs so that
    // two DMMA chains are in flight per item
    auto xt_item = [&](double* xt, const int item) {
        const int m = item / NNT, nt = item - m * NNT, j = 8 * m + g;
        double c0 = 0.0, c1 = 0.0, d0 = 0.0, d1 = 0.0;
        auto operands = [&](const int ks, double& av, double& bv) {   // (zero rows / columns beyond K)
            av = sT[(ks + t) * LDT + j];
            bv = sXM[(size_t)(ks + t) * LDX + 8 * nt + g];
        };
        int ks = 0;
#pragma unroll 2   // operands of the next pair of k-steps in flight while this pair's DMMA issue
        for (; ks + 4 < Kp4; ks += 8) {
            double a0, b0, a1, b1;
            operands(ks, a0, b0);
            operands(ks + 4, a1, b1);
            dmma8x8x4(c0, c1, a0, b0);
            dmma8x8x4(d0, d1, a1, b1);
        }
        if (ks < Kp4) {
            double a0, b0;
            operands(ks, a0, b0);
            dmma8x8x4(c0, c1, a0, b0);
        }
        *reinterpret_cast<double2*>(xt + (size_t)j * LD + 8 * nt + 2 * t) = make_double2(c0 + d0, c1 + d1);
    };
    // all TP / 8 path blocks of one row block at once, like a fit item: the T fragment is loaded once per k-step
    auto xt_row_item = [&](double* xt, const int m) {
        const double* trow = sT + t * LDT + 8 * m + g;
        const double* xcol = sXM + (size_t)t * LDX + g;
        double cf[NNT][2];
#pragma unroll
        for (int nt = 0; nt < NNT; nt++) cf[nt][0] = cf[nt][1] = 0.0;
#pragma unroll 2
        for (int ks = 0; ks < Kp4; ks += 4) {
            const double av = trow[ks * LDT];
#pragma unroll
            for (int nt = 0; nt < NNT; nt++) dmma8x8x4(cf[nt][0], cf[nt][1], av, xcol[(size_t)ks * LDX + 8 * nt]);
        }
#pragma unroll
        for (int nt = 0; nt < NNT; nt++) *reinterpret_cast<double2*>(xt + (size_t)(8 * m + g) * LD + 8 * nt + 2 * t) = make_double2(cf[nt][0], cf[nt][1]);
    };
    constexpr int kAuxFirst = 13;   // warps 13, 14 (workers without a decision item when the grid has <= 104 rows) and the producer
    const bool overlap_f0 = nbuf == 2 && NT <= kAuxFirst;

    int64_t tile = blockIdx.x;
    if (warp == kWorkers && tile < n_tiles) produce(tile, 0);

    for (int it = 0; tile < n_tiles; tile += gridDim.x, it++) {
        const int b = (nbuf == 2) ? (it & 1) : 0;
        const int64_t n0 = tile * TP;
        const double* vn = sVn + (size_t)b * GnP * LD;
        double* raw = sRaw + (size_t)b * L.raw_rows * LD;
        const double* sS = raw;   // row 0: spot of period p
        double* const sXT = sXT0 + ((overlap_f0 && (it & 1)) ? (size_t)8 * MT * LD : 0);   // x~ of this tile
        // development (LSMC_TRACE): phase timestamps of CTA 0, one row per warp and tile
        long long* tr = (q.trace && blockIdx.x == 0 && it < 64 && lane == 0) ? q.trace + ((size_t)warp * 64 + it) * 8 : nullptr;
        if (tr) tr[0] = clock64();

        // ================= phase F0: basis monomials of both periods from the raw rows =================
        if (any_copy) mbar_wait(&s_full[b], (uint32_t)((it / nbuf) & 1));   // this tile's rows have landed
        // the other buffer was last read in the previous tile's decision phase: the next tile's copies start now — a V_next tile of
        // a hundred 192-byte rows takes the TMA unit about 4 000 clocks — and have landed when the decision phase begins
        if (warp == kWorkers && nbuf == 2 && tile + gridDim.x < n_tiles) produce(tile + gridDim.x, b ^ 1);
        // (with two tile buffers the basis rows of every tile but a CTA's first were computed during the previous tile's decision
        //  phase, by the three warps that have no decision item — see phase D)
        if (!overlap_f0 || it == 0) basis_rows(b, n0, tid, (int)blockDim.x, [] { __syncthreads(); });

        if (tr) tr[1] = clock64();
        // ================= phase F1: preconditioned regressors + fitted continuation values =================
        // Work of this phase: NF fit items (8 next-grid rows x TP paths, TP / 8 DMMA per k-step) and NX x~ items (8 basis rows x
        // 8 paths, one DMMA per k-step), both on the FP64 tensor pipe.  Fit item f goes to warp f % 16; the warps left with one fit
        // item fewer take TP / 8 x~ items each first (the same number of DMMA), the rest is dealt round robin.
        const int NF = (q.mode == 0 && q.cont_mode == 0) ? (GnP >> 3) : 0;
        const int NX = q.accum_mode == 1 ? MT * NNT : 0;
        if (NX && (!overlap_f0 || it == 0)) {   // (with two tile buffers: done during the previous tile's decision phase)
            // the workers (the producer warp is busy issuing the next tile's copies) left with one fit item fewer take two whole row
            // blocks each; the rest goes round robin over the workers
            const int n_hi = NF ? NF - ((NF - 1) / kSweepWarps) * kSweepWarps : kSweepWarps;   // warps with the larger number of fit items
            const int n_lo = max(0, kWorkers - n_hi);
            const int n_rows = min(MT, 2 * n_lo);                                             // row blocks taken whole
            if (warp >= n_hi && warp < kWorkers)
                for (int m = warp - n_hi; m < n_rows; m += n_lo) xt_row_item(sXT, m);
            if (warp < kWorkers)
                for (int e = n_rows * NNT + warp; e < NX; e += kWorkers) xt_item(sXT, e);
        } else if (q.accum_mode == 2) {
            for (int path = tid; path < TP; path += blockDim.x) sXT[path] = (n0 + path < q.N) ? 1.0 : 0.0;
        }
        if (tr) tr[6] = clock64();
        if (q.mode == 0) {
            if (q.cont_mode == 0) {
                // one block of 8 next-grid rows per warp step, all TP / 8 path blocks at once: the coefficient fragment is loaded
                // once per k-step and feeds independent accumulator chains
                for (int mt = warp; mt < NF; mt += kSweepWarps) {
                    const double* brow = sBeta + (size_t)(8 * mt + g) * LDB + t;
                    const double* xcol = sXP + (size_t)t * LDX + g;
                    double cf[NNT][2];
#pragma unroll
                    for (int nt = 0; nt < NNT; nt++) cf[nt][0] = cf[nt][1] = 0.0;
#pragma unroll 2
                    for (int ks = 0; ks < Kp4; ks += 4) {
                        const double av = brow[ks];
#pragma unroll
                        for (int nt = 0; nt < NNT; nt++) dmma8x8x4(cf[nt][0], cf[nt][1], av, xcol[(size_t)ks * LDX + 8 * nt]);
                    }
                    double* crow = sC + (size_t)(8 * mt + g) * LD + 2 * t;
#pragma unroll
                    for (int nt = 0; nt < NNT; nt++) *reinterpret_cast<double2*>(crow + 8 * nt) = make_double2(cf[nt][0], cf[nt][1]);
                }
            } else {
                for (int i = tid; i < q.Gn * TP; i += blockDim.x) {
                    const int l = i / TP, c = i - l * TP;
                    sC[(size_t)l * LD + c] = __ldg(q.beta + (size_t)l * q.ldb);
                }
            }
        }
        if (tr) tr[2] = clock64();
        __syncthreads();
        if (tr) tr[3] = clock64();

        // ================= phase D: decisions + normal-equation accumulation; producer prefetches =================
        const int64_t next = tile + gridDim.x;
        if (accum) {
            // this warp's blocks of X~^T X~ (B operand = x~ itself); the slots' DMMA chains are independent
#pragma unroll
            for (int u = 0; u < TP / 8; u++) {
#pragma unroll
                for (int sl = 0; sl < GB; sl++) {
                    if (gblk[sl] >= 0) {
                        const double2 b2 = lds128(sXT + (size_t)(8 * (gblk[sl] >> 8) + g) * LD + 2 * t + 8 * u);
                        const double2 a2 = lds128(sXT + (size_t)(8 * (gblk[sl] & 255) + g) * LD + 2 * t + 8 * u);
                        dmma8x8x4(gacc[sl][0], gacc[sl][1], a2.x, b2.x);
                        dmma8x8x4(gacc[sl][0], gacc[sl][1], a2.y, b2.y);
                    }
                }
            }
        }
        if (overlap_f0 && warp >= kAuxFirst && next < n_tiles) {
            // the next tile's rows are on their way: as soon as they have landed, warps 13..15 run its basis program — the basis
            // rows, product nodes and the constant rows are not read during this phase
            if (any_copy) mbar_wait(&s_full[b ^ 1], (uint32_t)(((it + 1) >> 1) & 1));
            if (tr) tr[7] = clock64();
            basis_rows(b ^ 1, next * TP, tid - kAuxFirst * 32, (kSweepWarps - kAuxFirst) * 32, [] { asm volatile("bar.sync 1, 96;" ::: "memory"); });
            if (q.accum_mode == 1) {
                // ... and its preconditioned regressors, into the other x~ buffer: whole row blocks first, the rest as single items
                double* const xt_next = sXT0 + ((it & 1) ? 0 : (size_t)8 * MT * LD);
                constexpr int kAux = kSweepWarps - kAuxFirst, R = MT - MT % kAux;
                for (int m = warp - kAuxFirst; m < R; m += kAux) xt_row_item(xt_next, m);
                for (int e = R * NNT + warp - kAuxFirst; e < MT * NNT; e += kAux) xt_item(xt_next, e);
            }
        }
        if (warp < kWorkers) {
#pragma unroll
            for (int jj = 0; jj < NTW; jj++) {
                const int j = warp + jj * kWorkers;
                if (j >= n_items) break;
                {
                    const RowTable& R = tab[jj];
                    const int i = 8 * j + g;
                    double* vrow = q.Vout + (size_t)(R.row_ok ? i : 0) * q.Npad + n0;
                    if (q.mode == 1) terminal_item<TP, MT>(q, R, sS, sXT, vrow, accum, g, t, acc[jj]);
                    else switch (R.variant) {
#define VARIANT(f) case f: decide_item<TP, MT, f>(R, sC, vn, sS, sXT, vrow, q.df, accum, g, t, acc[jj]); break;
                        VARIANT(0) VARIANT(1) VARIANT(2) VARIANT(3) VARIANT(4) VARIANT(5) VARIANT(6) VARIANT(7)
                        VARIANT(kCons0 | kCons2 | kInvc)
#undef VARIANT
                        default: decide_item_generic<TP, MT>(q, R, i, sC, vn, sS, sXT, vrow, accum, g, t, acc[jj]); break;
                    }
                }
            }
        }
        if (tr) tr[4] = clock64();
        __syncthreads();
        if (tr) tr[5] = clock64();
        if (nbuf == 1) {
            if (warp == kWorkers && next < n_tiles) produce(next, 0);
        }
    }

    // ---- per-CTA partial: partial[cta][k][col], k = 8m+g, col = 8j+2t ----
    if (q.accum_mode) {
        // X~^T X~: block (kb, m) holds rows 8m.. of the columns of basis block kb; its mirror image is left to the solver
        // (SolveParams::gram_lower_blocks) and written as zeros so that the reduction never sums stale memory
        double* out = q.partial + (size_t)blockIdx.x * (8 * MT) * q.ncols;
#pragma unroll
        for (int sl = 0; sl < GB; sl++) {
            if (gblk[sl] < 0) continue;
            const int kb = gblk[sl] >> 8, m = gblk[sl] & 255;
            *reinterpret_cast<double2*>(out + (size_t)(8 * m + g) * q.ncols + 8 * (NT + kb) + 2 * t) = make_double2(gacc[sl][0], gacc[sl][1]);
            if (kb != m) *reinterpret_cast<double2*>(out + (size_t)(8 * kb + g) * q.ncols + 8 * (NT + m) + 2 * t) = make_double2(0.0, 0.0);
        }
    }
    if (q.accum_mode && warp < kWorkers) {
        double* out = q.partial + (size_t)blockIdx.x * (8 * MT) * q.ncols;
#pragma unroll
        for (int jj = 0; jj < NTW; jj++) {
            const int j = warp + jj * kWorkers;
            if (j >= n_items) break;
#pragma unroll
            for (int m = 0; m < MT; m++)
                *reinterpret_cast<double2*>(out + (size_t)(8 * m + g) * q.ncols + 8 * j + 2 * t) = make_double2(acc[jj][m][0], acc[jj][m][1]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Launch configuration
// ------------------------------------------------------------------------------------------------
static size_t layout_bytes(int TP, int Gn, int K, int F, int MT, int nbuf, int n_scratch, int n_rowops) {
    return sweep_layout(TP, Gn, K, F, MT, nbuf, n_scratch, n_rowops).total * sizeof(double);
}

// Largest tile that fits the 227 KB of an SM, double-buffered when possible.
bool choose_sweep_shape(int max_grid, int K, int F, int MT, int n_scratch, int n_rowops, int* TP, int* nbuf) {
    const size_t limit = 224 * 1024;
    const int tps[2] = {56, 24};
    for (int nb = 2; nb >= 1; nb--)
        for (int tp : tps)
            if (layout_bytes(tp, max_grid, K, F, MT, nb, n_scratch, n_rowops) <= limit) { *TP = tp; *nbuf = nb; return true; }
    return false;
}

template <int TP, int MT, int NTW>
static cudaError_t launch_sweep_t(const SweepParams& q, int blocks, size_t smem, int nbuf, cudaStream_t stream) {
    auto kern = sweep_kernel<TP, MT, NTW>;
    if (cudaError_t e = ensure_dynamic_smem((const void*)kern, smem)) return e;
    CUtensorMap mv{};
    if (q.mode == 0 && !make_panel_map(&mv, q.Vnext, q.Npad, q.Gn, TP, q.Gn)) return cudaErrorInvalidValue;
    return launch_pdl(q.pdl != 0, kern, blocks, kSweepWarps * 32, smem, stream, q, nbuf, mv);
}

template <int TP>
static cudaError_t launch_sweep_tp(const SweepParams& q, int MT, int NTW, int blocks, size_t smem, int nbuf, cudaStream_t s) {
#define ROW(mt) \
    if (MT == mt) { \
        if (NTW == 1) return launch_sweep_t<TP, mt, 1>(q, blocks, smem, nbuf, s); \
        return launch_sweep_t<TP, mt, 2>(q, blocks, smem, nbuf, s); \
    }
#ifdef LSMC_FAST_BUILD   // development builds: only the shapes of configs C1-C3/C5
    ROW(1) ROW(2)
#else
    ROW(1) ROW(2) ROW(4) ROW(7) ROW(8)
#endif
#undef ROW
    return cudaErrorInvalidValue;
}

int sweep_max_items() { return 2 * kWorkers; }

cudaError_t launch_sweep(SweepParams& q, const SweepConfig& cfg, cudaStream_t stream) {
    const int MT = cfg.MT;
    const int NT = (q.G + 7) / 8;
    if (NT + (q.accum_mode ? MT : 0) > 2 * kWorkers) return cudaErrorInvalidValue;   // (the bound lsmc_set_problem checks: sweep_max_items)
    const int NTW = NT <= kWorkers ? 1 : 2;
    q.ncols = 8 * (NT + MT);
    const size_t smem = layout_bytes(cfg.TP, q.Gn, q.K, q.F, MT, cfg.nbuf, cfg.n_scratch, cfg.n_rowops);
    if (cfg.TP == 56) return launch_sweep_tp<56>(q, MT, NTW, cfg.blocks, smem, cfg.nbuf, stream);
#ifdef LSMC_FAST_BUILD
    return cudaErrorInvalidValue;
#else
    return launch_sweep_tp<24>(q, MT, NTW, cfg.blocks, smem, cfg.nbuf, stream);
#endif
}

}  // namespace lsmc
