// forward2.cu — K4, streaming form of the straight-line forward kernel (kernels.cu forward_fast_kernel; LSV.cs:393-480).
//
// Same arithmetic, bit for bit (tests/test_gpu_parity.py::test_forward_straight_line_kernel_equals_the_bisecting_kernel),
// reorganised around what the profile of the old kernel showed (profiles/r01_forward_simulate_ncu_summary_s21c.txt: 0.06
// of the HBM peak, 230 M local loads, IPC 0.41 at 16 warps per SM, two CTA barriers + a block reduction + a chain of
// dependent global loads per period):
//   * the constants of a period — ForwardParams, the coefficient rows beta[Gn][K], the next grid, the ratchet pillars with
//     their slopes — are a pure function of the period index, so they are copied into a double-buffered shared-memory
//     stage with cp.async while the previous period is being computed: one CTA barrier per period, no exposed load;
//   * per-period sums leave the kernel per WARP (shuffle tree, one partial row per warp in global memory): no shared-memory
//     reduction, no second barrier;
//   * coefficient rows at an even pitch with odd half (K = 10 -> 10): 128-bit loads, two fused multiply-adds per load;
//   * 32-bit shared-memory indexing.
// Measured (1M paths, C3): 27.8 -> 24.6 ms; 125k paths: 7.9 -> 4.9 ms (the per-period fixed cost is what shrank).  The path
// body itself is unchanged: ~1 100 warp instructions per path-step of which ~230 are FP64 arithmetic (ncu: issue slots 44 %
// busy, FP64 pipe 24 %, stalls short-scoreboard 2.4 / wait 2.4 / long-scoreboard 1.4 per issue at 16 warps per SM; three CTAs
// per SM at 85 registers spill and are slower: 31 ms).
#include "kernels.h"

namespace lsmc {

namespace {

constexpr int kF2Threads = 256;
constexpr int kF2Warps = kF2Threads / 32;
#ifndef LSMC_F2_MINB
#define LSMC_F2_MINB 2
#endif
constexpr int kF2MinBlocks = LSMC_F2_MINB;
constexpr int kF2MaxGrid = 127;

__device__ __forceinline__ void cp_async8(void* dst_smem, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Result of the bisection of SH.cs:280-314 per comparison class: see kernels.cu fwd_class_result (same function; kept in
// both translation units because it is inlined into the table builders).
__device__ __forceinline__ uint16_t f2_class_result(int n, int m, int flags) {
    const int top = n - 1;
    auto near = [&](int x) {
        return (x == top - 1 && (flags & 4)) || (x == top && (flags & 8)) || (x == m && (flags & 1)) || (x == m + 1 && (flags & 2));
    };
    if (n == 1) return near(0) ? (uint16_t)0 : (uint16_t)0xffff;
    int left = 0, right = top;
    for (int guard = 0; guard < 64 && right > left; guard++) {
        const int mid = (left + right) >> 1;
        if (near(mid)) return (uint16_t)(mid | (mid << 8));
        if (mid >= m + 1) { right = mid; continue; }                                   // grid[mid] > t
        if (mid + 1 >= m + 1) return (uint16_t)(mid | ((mid + 1) << 8));               // t <= grid[mid + 1]
        if (near(mid + 1)) return (uint16_t)((mid + 1) | ((mid + 1) << 8));
        if (mid + 1 == top) return (uint16_t)0xffff;
        left = mid;
    }
    return (uint16_t)0xffff;
}

// One period's constants in shared memory.  Offsets in doubles from the stage base.
struct F2Layout {
    int BP, o_beta, o_grid, o_pill, o_q, stage, o_nodes, o_stage0, o_cls, total_doubles;
};
__host__ __device__ inline F2Layout f2_layout(int K, int n_nodes, int max_grid) {
    F2Layout L;
    L.BP = forward2_beta_pitch(K);
    L.o_beta = 0;
    L.o_grid = L.o_beta + max_grid * L.BP;
    L.o_pill = L.o_grid + ((max_grid + 1) & ~1);
    L.o_q = L.o_pill + kFwd2PillarDoubles;
    L.stage = L.o_q + (int)((sizeof(ForwardParams) + 15) / 16) * 2;
    L.o_nodes = 0;
    L.o_stage0 = n_nodes * kF2Threads;
    L.o_cls = L.o_stage0 + 2 * L.stage;
    L.total_doubles = L.o_cls + ((kF2MaxGrid + 1) * 16 * 2 + 7) / 8;
    return L;
}

}  // namespace

// KT > 0: K = KT at compile time (K <= 16): the path's basis row and its column sums are registers.
// KT == 0: any K <= 64 (C4: 56 monomials of five factors).  The basis row stays in the thread's shared-memory node column, the
//          six fitted values of a path come from one pass over k, and the design matrix's column sums are taken by the warp
//          together from its 32 node columns (lane = basis function, rotated reads: conflict free) — 112 accumulator registers per
//          thread was what made the general forward kernel spill (342 ms per valuation at C4).
template <int KT>
__global__ void __launch_bounds__(kF2Threads, KT == 0 ? 1 : kF2MinBlocks) forward2_kernel(const Forward2Args a) {
    extern __shared__ __align__(16) double f2sm[];
    constexpr bool BIG = KT == 0;
    constexpr int KR = BIG ? 1 : KT;           // register arrays of the small-K form
    const int K = BIG ? a.K : KT, nsum = kFwdSums + K;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const F2Layout L = f2_layout(K, a.n_nodes, a.max_grid);
    const int BP = L.BP;
    double* const nodes = f2sm + tid;                       // node j of this thread at nodes[j * kF2Threads]
    uint16_t* const s_class = reinterpret_cast<uint16_t*>(f2sm + L.o_cls);
    __shared__ uint4 s_ops[kMaxOps];                        // basis program as element offsets into the node column: dst, a, b
    __shared__ uint32_t s_basis_off[64];
    {
        const BasisDev* __restrict__ bd0 = a.basis;
        for (int o = tid; o < bd0->n_ops; o += kF2Threads) {
            const BasisOpDev op = bd0->ops[o];
            s_ops[o] = make_uint4((uint32_t)op.dst * kF2Threads, (uint32_t)op.a * kF2Threads, (uint32_t)op.b * kF2Threads, 0u);
        }
        for (int k = tid; k < 64; k += kF2Threads) s_basis_off[k] = (uint32_t)bd0->node_of_basis[k] * kF2Threads;
    }
    const int n_ops = a.basis->n_ops, F = a.F;

    // asynchronous fill of a stage with the constants of period p: addresses are functions of p alone
    auto fill = [&](int p, int buf) {
        double* st = f2sm + L.o_stage0 + buf * L.stage;
        const double* beta = a.beta_all + (size_t)p * a.beta_stride;
        const int per_row = BP / 2;                          // 16-byte chunks per staged row (ldb >= BP, rows 16-byte aligned)
        for (int e = tid; e < a.max_grid * per_row; e += kF2Threads) {
            const int l = e / per_row, c = e - l * per_row;
            cp_async16(st + L.o_beta + l * BP + 2 * c, beta + (size_t)l * a.ldb + 2 * c);
        }
        const double* grid = a.grid_pad + (size_t)(p + 1) * a.grid_pitch;
        for (int e = tid; e < a.grid_pitch / 2; e += kF2Threads) cp_async16(st + L.o_grid + 2 * e, grid + 2 * e);
        const double* pill = a.pillars + (size_t)p * kFwd2PillarDoubles;
        for (int e = tid; e < kFwd2PillarDoubles / 2; e += kF2Threads) cp_async16(st + L.o_pill + 2 * e, pill + 2 * e);
        const double* qsrc = reinterpret_cast<const double*>(a.params + p);
        for (int e = tid; e < (int)(sizeof(ForwardParams) / 8); e += kF2Threads) cp_async8(st + L.o_q + e, qsrc + e);
    };

    fill(a.p0, a.p0 & 1);
    int class_n = -1;
    const unsigned warp_global = blockIdx.x * kF2Warps + warp;
    const int64_t stride = (int64_t)gridDim.x * kF2Threads;

    for (int p = a.p0; p < a.p1; p++) {
        cp_async_wait_all();
        __syncthreads();   // stage p is complete, and everybody has left stage p-1 (= the buffer filled next)
        if (p + 1 < a.p1) fill(p + 1, (p + 1) & 1);
        const double* st = f2sm + L.o_stage0 + (p & 1) * L.stage;
        const ForwardParams& q = *reinterpret_cast<const ForwardParams*>(st + L.o_q);
        const double* beta_s = st + L.o_beta;
        const double* grid_s = st + L.o_grid;
        const double* s_cx = st + L.o_pill;
        const double* s_clo = s_cx + 16; const double* s_chi = s_cx + 32; const double* s_slo = s_cx + 48; const double* s_shi = s_cx + 64;
        const int Gn = q.Gn, top = Gn - 1;
        if (Gn != class_n) {   // the table is a function of the number of grid points alone (ramp-up / ramp-down periods)
            __syncthreads();   // nobody still reads the old table (they all passed the barrier above after their last path)
            for (int e = tid; e < (Gn + 1) * 16; e += kF2Threads) s_class[e] = f2_class_result(Gn, (e >> 4) - 1, e & 15);
            class_n = Gn;
            __syncthreads();
        }
        const bool fitted_mode = q.cont_mode == 0;
        const double g0 = grid_s[0], inv_h = Gn > 1 ? 1.0 / (grid_s[1] - grid_s[0]) : 0.0;
        const double g_top = grid_s[top], g_pre = grid_s[max(top - 1, 0)];
        const double df = q.df, tol = q.tol, loss_pct = q.loss_pct, next_min = q.next_min, next_max = q.next_max;
        const bool spot_is_const = q.spot_is_const != 0;
        const double* __restrict__ spot_row = q.spot;
        const double* __restrict__ fac_rows = q.fac;
        double* __restrict__ inv_state = q.inventory;
        double* __restrict__ pv_state = q.pv_by_sim;
        const int64_t N = q.N, Npad = q.Npad;
        const int c_kind = q.c_kind, c_n = q.c_n;
        const bool any_panel = q.o_inventory || q.o_volume || q.o_consumed || q.o_loss || q.o_net || q.o_pv || q.o_decision;

        double sums[kFwdSums], xsum[KR];
        double xs0 = 0.0, xs1 = 0.0;   // BIG: column sums of basis functions `lane` and `lane + 32` over the warp's paths
#pragma unroll
        for (int i = 0; i < kFwdSums; i++) sums[i] = 0.0;
#pragma unroll
        for (int k = 0; k < KR; k++) xsum[k] = 0.0;

        // (BIG: the path loop is warp uniform — the warp sums its node columns together — and lanes beyond N idle inside it)
        for (int64_t nb = (int64_t)blockIdx.x * kF2Threads + (BIG ? warp * 32 : tid); nb < N; nb += stride) {
            const int64_t n = BIG ? nb + lane : nb;
            const bool live = !BIG || n < N;
            if (BIG && fitted_mode) {
                if (live) {
                    nodes[0] = 0.0; nodes[kF2Threads] = 1.0; nodes[2 * kF2Threads] = spot_is_const ? q.spot_const : __ldcs(spot_row + n);
                    const double* fp = fac_rows + n;
                    for (int f = 0; f < F; f++, fp += Npad) nodes[(3 + f) * kF2Threads] = __ldcs(fp);
                    for (int o = 0; o < n_ops; o++) {
                        const uint4 op = s_ops[o];
                        nodes[op.x] = mul_rn(nodes[op.y], nodes[op.z]);
                    }
                } else {
                    for (int j = 0; j < a.n_nodes; j++) nodes[j * kF2Threads] = 0.0;   // a lane beyond the last path adds zeros to the column sums
                }
                __syncwarp();
                {
                    const double* col = f2sm + warp * 32;
                    const uint32_t o0 = s_basis_off[lane], o1 = s_basis_off[lane + 32];
#pragma unroll 8
                    for (int t = 0; t < 32; t++) {
                        const int r = (t + lane) & 31;
                        if (lane < K) xs0 += col[o0 + r];
                        if (lane + 32 < K) xs1 += col[o1 + r];
                    }
                }
                __syncwarp();
            }
            if (!live) continue;
            const double s = spot_is_const ? q.spot_const : __ldcs(spot_row + n);
            const double v = inv_state[n];
            const double pv_old = pv_state[n];
            double xrow[KR];
            if (BIG) {
                // (basis row already in the node column)
            } else if (fitted_mode) {
                nodes[0] = 0.0; nodes[kF2Threads] = 1.0; nodes[2 * kF2Threads] = s;
                {
                    const double* fp = fac_rows + n;
                    for (int f = 0; f < F; f++, fp += Npad) nodes[(3 + f) * kF2Threads] = __ldcs(fp);
                    if (n + stride < N) {   // the next path's rows on their way to L2 while this one is worked on
                        const double* nx = fac_rows + n + stride;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(spot_row + n + stride));
                        for (int f = 0; f < F; f++, nx += Npad) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx));
                    }
                }
                for (int o = 0; o < n_ops; o++) {
                    const uint4 op = s_ops[o];
                    nodes[op.x] = mul_rn(nodes[op.y], nodes[op.z]);
                }
#pragma unroll
                for (int k = 0; k < KT; k++) { xrow[k] = nodes[s_basis_off[k]]; xsum[k] += xrow[k]; }
            } else {
#pragma unroll
                for (int k = 0; k < KT; k++) xrow[k] = k == 0 ? 1.0 : 0.0;
            }
            // CmdtyStorage.GetInjectWithdrawRange (CmdtyStorage.cs:86-100)
            if (v < q.min_inv - q.slack_lo || v > q.max_inv + q.slack_hi) atomicExch(q.error_flag, 1);
            double rmin, rmax;
            if (c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {   // LinearSpline: segment by counting, slopes precomputed on the host
                int k = 0;
#pragma unroll
                for (int i = 1; i <= 6; i++) k += (s_cx[i] <= v) ? 1 : 0;   // entries from c_n - 1 on are +inf
                for (int i = 7; i < c_n - 1; i++) k += (s_cx[i] <= v) ? 1 : 0;
                const double dxv = sub_rn(v, s_cx[k]);
                rmax = add_rn(s_chi[k], mul_rn(dxv, s_shi[k]));
                rmin = add_rn(s_clo[k], mul_rn(dxv, s_slo[k]));
            } else if (c_kind == LSMC_CONSTRAINT_CONSTANT) { rmin = s_clo[0]; rmax = s_chi[0]; }
            else if (c_kind == LSMC_CONSTRAINT_POLYNOMIAL) {   // Polynomial.Evaluate: Horner over the staged coefficients, no contraction
                rmax = s_chi[c_n - 1]; rmin = s_clo[c_n - 1];
                for (int i = c_n - 2; i >= 0; i--) { rmax = add_rn(mul_rn(rmax, v), s_chi[i]); rmin = add_rn(mul_rn(rmin, v), s_clo[i]); }
            } else {   // step ratchets (StepInjectWithdrawConstraint.cs:72-79): the pillar at or below v; entries from c_n on are +inf
                if (v < s_cx[0] - q.slack_hi || v > s_cx[c_n - 1] + q.slack_hi) atomicExch(q.error_flag, 2);
                int idx = 0;
                for (int i = 1; i < c_n; i++) idx += (s_cx[i] <= v) ? 1 : 0;
                rmin = s_clo[idx]; rmax = s_chi[idx];
            }
            const double loss = mul_rn(loss_pct, v);
            // bang-bang decision set (SH.cs:109-197), extra decisions == 0
            double w, jn;
            {
                const double base = sub_rn(v, loss);
                const double afterW = add_rn(rmin, base), afterJ = add_rn(rmax, base);
                if (afterW > next_max && !(sub_rn(afterW, next_max) < tol)) atomicExch(q.error_flag, 3);
                if (afterJ < next_min && !(sub_rn(next_min, afterJ) < tol)) atomicExch(q.error_flag, 3);
                w = afterW > next_max ? sub_rn(next_max, base) : (afterW > next_min ? rmin : sub_rn(next_min, base));
                jn = afterJ < next_min ? sub_rn(next_min, base) : (afterJ < next_max ? rmax : sub_rn(next_max, base));
            }
            const bool one_sided = (w >= 0.0 || jn <= 0.0);
            const double inv_cost_npv = q.inv_cost == 0.0 ? 0.0 : mul_rn(mul_rn(v, q.inv_cost), q.inv_cost_df);
            double vol[3], tot[3], imm[3], cons[3];
            vol[0] = w; vol[1] = one_sided ? jn : 0.0; vol[2] = jn;
            bool all_found = true;
            double flo[3], fhi[3], gl_d[3], gu_d[3], after_d[3];
            bool lo_eq_hi[3];
            [[maybe_unused]] int blo[3], bhi[3];
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double vd = vol[d];
                const double after = sub_rn(add_rn(v, vd), loss);
                const bool pos = vd > 0.0;
                const double avol = fabs(vd);
                cons[d] = mul_rn(pos ? q.cons_inj : q.cons_wd, avol);
                const double ia = mul_rn(mul_rn(-vd, s), df);
                const double ic = mul_rn(mul_rn(-cons[d], s), df);
                const double cost = mul_rn(mul_rn(pos ? q.inj_cost : q.wd_cost, avol), pos ? q.inj_df : q.wd_df);
                imm[d] = sub_rn(add_rn(sub_rn(ia, cost), ic), inv_cost_npv);
                // the bisection's result in O(1): interval by guess + one correction, comparison class, class table
                int m = (int)((after - g0) * inv_h);
                m = min(max(m, 0), max(top - 1, 0));
                const double ga = grid_s[max(m - 1, 0)], gb = grid_s[m], gc = grid_s[min(m + 1, top)], gd = grid_s[min(m + 2, top)];
                const bool down = after <= gb, up = !down && after > gc;
                m = down ? m - 1 : (up ? m + 1 : m);
                double gl = down ? ga : (up ? gc : gb);
                double gu = down ? gb : (up ? gd : gc);
                const bool below = after <= g0, above = after > g_top;
                m = below ? -1 : (above ? top : m);
                gl = above ? g_top : gl;
                gu = below ? g0 : gu;
                const int flags = ((!below && fabs(after - gl) <= tol) ? 1 : 0) | ((!above && fabs(after - gu) <= tol) ? 2 : 0) |
                                  ((top >= 1 && fabs(after - g_pre) <= tol) ? 4 : 0) | ((fabs(after - g_top) <= tol) ? 8 : 0);
                const unsigned cls = s_class[(m + 1) * 16 + flags];
                const bool found = cls != 0xffffu;
                const int lo = found ? (int)(cls & 0xffu) : 0, hi = found ? (int)(cls >> 8) : 0;
                gl = grid_s[lo]; gu = grid_s[hi];
                all_found = all_found && found;
                double f_lo = 0.0, f_hi = 0.0;
                if (BIG) { blo[d] = lo * BP; bhi[d] = hi * BP; }
                else {
                    const double2* bl = reinterpret_cast<const double2*>(beta_s + lo * BP);
                    const double2* bh = reinterpret_cast<const double2*>(beta_s + hi * BP);
#pragma unroll
                    for (int k2 = 0; k2 < (KR + 1) / 2; k2++) {
                        const double2 cl = bl[k2], ch = bh[k2];
                        f_lo = fma(xrow[2 * k2], cl.x, f_lo); f_hi = fma(xrow[2 * k2], ch.x, f_hi);
                        if (2 * k2 + 1 < KR) { f_lo = fma(xrow[2 * k2 + 1], cl.y, f_lo); f_hi = fma(xrow[2 * k2 + 1], ch.y, f_hi); }
                    }
                }
                flo[d] = f_lo; fhi[d] = f_hi; lo_eq_hi[d] = lo == hi; gl_d[d] = gl; gu_d[d] = gu; after_d[d] = after;
            }
            if (BIG) {   // the six fitted values in one pass over k: x_k from the node column, coefficient pairs from the staged rows
                if (fitted_mode) {
                    for (int k2 = 0; k2 < (K + 1) / 2; k2++) {
                        const double x0 = nodes[s_basis_off[2 * k2]];
                        const double x1 = (2 * k2 + 1 < K) ? nodes[s_basis_off[2 * k2 + 1]] : 0.0;
#pragma unroll
                        for (int d = 0; d < 3; d++) {
                            const double2 cl = *reinterpret_cast<const double2*>(beta_s + blo[d] + 2 * k2);
                            const double2 ch = *reinterpret_cast<const double2*>(beta_s + bhi[d] + 2 * k2);
                            flo[d] = fma(x0, cl.x, flo[d]); fhi[d] = fma(x0, ch.x, fhi[d]);
                            if (2 * k2 + 1 < K) { flo[d] = fma(x1, cl.y, flo[d]); fhi[d] = fma(x1, ch.y, fhi[d]); }
                        }
                    }
                } else {
#pragma unroll
                    for (int d = 0; d < 3; d++) { flo[d] = beta_s[blo[d]]; fhi[d] = beta_s[bhi[d]]; }   // x = (1, 0, ...)
                }
            }
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double f_lo = flo[d], f_hi = fhi[d], gl = gl_d[d], gu = gu_d[d], after = after_d[d];
                const bool same = lo_eq_hi[d];
                const double wl = sub_rn(gu, after) / sub_rn(gu, gl);
                const double wu = sub_rn(1.0, wl);
                const double cont = same ? f_lo : add_rn(mul_rn(f_lo, wl), mul_rn(f_hi, wu));
                tot[d] = add_rn(imm[d], cont);
            }
            if (!all_found) atomicExch(q.error_flag, 4);
            // first strictly greatest (SH.cs:206-222)
            const bool t1 = tot[1] > tot[0];
            double best = t1 ? tot[1] : tot[0];
            double best_imm = t1 ? imm[1] : imm[0], best_cons = t1 ? cons[1] : cons[0], best_vol = t1 ? vol[1] : vol[0];
            int best_d = t1 ? 1 : 0;
            const bool t2 = tot[2] > best;
            best_imm = t2 ? imm[2] : best_imm; best_cons = t2 ? cons[2] : best_cons; best_vol = t2 ? vol[2] : best_vol;
            best_d = t2 ? 2 : best_d;
            const double next_inv = sub_rn(add_rn(v, best_vol), loss);
            inv_state[n] = next_inv;
            pv_state[n] = add_rn(pv_old, best_imm);
            const double net = sub_rn(-best_vol, best_cons);
            sums[0] += v; sums[1] += best_vol; sums[2] += best_cons; sums[3] += loss; sums[4] += net; sums[5] += best_imm;
            sums[6] += mul_rn(-(add_rn(best_vol, best_cons)), s);
            if (any_panel) {
                if (q.o_inventory) __stcs(q.o_inventory + n, v);
                if (q.o_volume) __stcs(q.o_volume + n, best_vol);
                if (q.o_consumed) __stcs(q.o_consumed + n, best_cons);
                if (q.o_loss) __stcs(q.o_loss + n, loss);
                if (q.o_net) __stcs(q.o_net + n, net);
                if (q.o_pv) __stcs(q.o_pv + n, best_imm);
                if (q.o_decision) q.o_decision[n] = (int8_t)best_d;
            }
        }
        // per-warp partial sums (fixed shuffle tree), one row per warp: sums[p][warp_global][nsum]
        double* out = a.warp_sums + ((size_t)p * a.n_parts + warp_global) * nsum;
#pragma unroll
        for (int i = 0; i < kFwdSums + KR; i++) {
            if (BIG && i >= kFwdSums) break;
            double val = i < kFwdSums ? sums[i < kFwdSums ? i : 0] : xsum[i >= kFwdSums ? i - kFwdSums : 0];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) val += __shfl_xor_sync(0xffffffffu, val, off);
            if (lane == 0) out[i] = val;
        }
        if (BIG) {   // the column sums are already whole-warp sums: lane = basis function
            if (lane < K) out[kFwdSums + lane] = xs0;
            if (lane + 32 < K) out[kFwdSums + lane + 32] = xs1;
        }
    }
}

int forward2_blocks_per_sm(int K) { return K > 16 ? 1 : kF2MinBlocks; }
int forward2_threads() { return kF2Threads; }
int forward2_parts(int blocks) { return blocks * kF2Warps; }

size_t forward2_smem_bytes(int K, int n_nodes, int max_grid) { return sizeof(double) * (size_t)f2_layout(K, n_nodes, max_grid).total_doubles; }

template <int KT>
static cudaError_t launch_forward2_t(const Forward2Args& a, int blocks, cudaStream_t stream) {
    const size_t smem = forward2_smem_bytes(KT, a.n_nodes, a.max_grid);
    if (cudaError_t e = ensure_dynamic_smem((const void*)forward2_kernel<KT>, smem)) return e;
    forward2_kernel<KT><<<blocks, kF2Threads, smem, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_forward2(const Forward2Args& a, int K, int blocks, cudaStream_t stream) {
    if (K > 16) {
        if (K > 64 || a.K != K) return cudaErrorInvalidValue;
        const size_t smem = forward2_smem_bytes(K, a.n_nodes, a.max_grid);
        if (cudaError_t e = ensure_dynamic_smem((const void*)forward2_kernel<0>, smem)) return e;
        forward2_kernel<0><<<blocks, kF2Threads, smem, stream>>>(a);
        return cudaGetLastError();
    }
    switch (K) {
#define F2_K(k) case k: return launch_forward2_t<k>(a, blocks, stream);
#ifdef LSMC_FAST_BUILD
        F2_K(3) F2_K(10)
#else
        F2_K(1) F2_K(2) F2_K(3) F2_K(4) F2_K(5) F2_K(6) F2_K(7) F2_K(8) F2_K(9) F2_K(10)
        F2_K(11) F2_K(12) F2_K(13) F2_K(14) F2_K(15) F2_K(16)
#endif
#undef F2_K
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace lsmc
