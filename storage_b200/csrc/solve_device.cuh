// solve_device.cuh — K2b as a CTA-level routine: the normal equations of one period for a set of grid points.
//   red(k, col): col < G -> X~^T V_col ; col = 8*NT + k' -> X~^T X~  (the reduced, all-reduced normal-equation block).
// Column-equilibrated Cholesky (G~ = D G D, D = diag(G)^-1/2) with a pivot guard that drops numerically dependent
// columns, one step of iterative refinement against the unscaled system, beta = T (D x).
// Used by solve_kernel (one CTA, all right-hand sides; kernels.cu) and by tail_kernel (tail.cu: every CTA factorises
// redundantly and takes the right-hand sides col0, col0 + col_step, ...).
#pragma once
#include "kernels.h"

namespace lsmc {

// The reduced block in one device buffer.
struct GlobalRed {
    const double* __restrict__ p;
    int ld;
    __device__ __forceinline__ double operator()(int k, int col) const { return p[(size_t)k * ld + col]; }
};

// Shared memory of a solve, in doubles: A [K][K] (Cholesky factor, lower) | Graw [K][K] | dscale [K] | Tm [K][K] | invd [K] |
// S [K][K] (the factor in the form the warp substitutions read, see solve_factor) | col [2][K] (published columns).
inline size_t solve_smem_bytes(int K) { return sizeof(double) * ((size_t)4 * K * K + 4 * K); }

// Element idx of the lower triangle in row-major order (0: (0,0), 1: (1,0), 2: (1,1), ...) -> row; column = idx - row (row + 1) / 2.
static __device__ __noinline__ int tri_row(int idx) {   // (one copy: the tail kernel runs its code once per launch, from a cold instruction cache)
    int i = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
    while ((i + 1) * (i + 2) / 2 <= idx) i++;
    while (i * (i + 1) / 2 > idx) i--;
    return i;
}

// The preconditioner T (upper triangular: only that triangle is read) -> shared memory.  Per-problem constant data: tail_kernel
// calls this before its dependency wait, so the loads overlap the drain of the sweep before it.
__device__ __forceinline__ void solve_stage_precond(const SolveParams& sp, double* sm) {
    const int K = sp.K;
    const double* T = sp.precond;
    if (sp.accum_mode == 2 || !T) return;
    double* Tm = sm + 2 * K * K + K;
    const int n_tri = K * (K + 1) / 2;
#pragma unroll 8
    for (int e = threadIdx.x; e < n_tri; e += blockDim.x) {
        const int m = tri_row(e), k = e - m * (m + 1) / 2;   // k <= m
        Tm[k * K + m] = __ldg(T + k * K + m);
    }
}

// Gram matrix -> shared memory, equilibration, preconditioner.  Returns false in the current-period mode (means, no system).
// Leaves Graw, dscale, dropped (and Tm unless the caller staged it already); the scaled matrix D G D is formed by solve_factor
// in registers.
template <class Red>
__device__ __forceinline__ bool solve_stage(const SolveParams& sp, const Red& red, double* sm, int* dropped, const int col0, const int col_step,
                                            const bool precond_staged = false) {
    const int K = sp.K, G = sp.G, tid = threadIdx.x;
    if (sp.accum_mode == 2) {   // current period: continuation = mean over sims (LSV.cs:171-182)
        for (int i = col0 + tid * col_step; i < G; i += blockDim.x * col_step) {
            double* b = sp.beta + (size_t)i * sp.ldb;
            b[0] = red(0, i) / (double)sp.n_global;
            for (int k = 1; k < sp.ldb; k++) b[k] = 0.0;
        }
        return false;
    }
    double* A = sm;                 // [K][K] Cholesky factor (lower), written by solve_factor
    double* Graw = A + K * K;       // [K][K] unscaled Gram
    double* dscale = Graw + K * K;  // [K]
    const int gcol = 8 * sp.NT;
    const double* T = sp.precond;
    const int n_tri = K * (K + 1) / 2;
    // symmetrise: the two triangles come from different DMMA tiles and may differ in the last bit.  One pair of loads per
    // element of the lower triangle, eight elements (sixteen L2 loads) in flight per thread.
#pragma unroll 8
    for (int e = tid; e < n_tri; e += blockDim.x) {
        const int r = tri_row(e), c = e - r * (r + 1) / 2;
        // (c <= r; with gram_lower_blocks only the blocks with basis block <= column block are filled and their mirror images hold
        //  zeros: (c, r) then carries the whole entry.  Both loads unconditional: sixteen in flight per thread)
        const double w = (sp.gram_lower_blocks && (r >> 3) != (c >> 3)) ? 1.0 : 0.5;
        const double v = w * (red(r, gcol + c) + red(c, gcol + r));
        Graw[r * K + c] = v;
        Graw[c * K + r] = v;
    }
    if (!precond_staged) solve_stage_precond(sp, sm);
    __syncthreads();
    for (int k = tid; k < K; k += blockDim.x) {
        const double gkk = Graw[k * K + k];
        const bool flagged = T && T[(size_t)K * K + k] != 0.0;   // column found dependent by the preconditioner's QR
        dscale[k] = (gkk > 0.0 && !flagged) ? rsqrt(gkk) : 0.0;
        dropped[k] = (gkk > 0.0 && !flagged) ? 0 : 1;
    }
    __syncthreads();
    return true;
}

// Cholesky of the staged Gram matrix, equilibrated on the fly, by the whole CTA + reciprocals of the diagonal (CTA-wide call).
// Right-looking with the scaling deferred:
//     a[i][k] -= (a[i][j] / pivot_j) a[k][j]          (j < k <= i)
// Every thread keeps its E elements of the lower triangle (dealt row-major, round robin) in REGISTERS for the whole
// factorisation; at column j the owners of that column publish it — it is final — into one of two K-vectors in shared memory,
// one barrier, and every thread updates its elements right of the column from two shared loads each.  L = a scaled by
// 1 / sqrt(pivot) is written once at the end, as A (lower triangle, with its diagonal) for the thread-per-right-hand-side
// solver and as S for the warp solver: S[j][k] = l_jk / l_jj left of the diagonal (row j of the backward substitution) and
// S[j][k] = l_kj / l_jj right of it (column j of the forward substitution), so that both substitutions read ROW j — lanes on
// consecutive addresses.  Same operations on the same values as the shared-memory form it replaces (which walked the trailing
// triangle with two nested strided loops: 0.83 us per column at K = 56), element by element.
template <int E>
__device__ __forceinline__ void solve_factor(const SolveParams& sp, double* sm, int* dropped) {
    const int K = sp.K, tid = threadIdx.x, nthr = blockDim.x;
    double* A = sm;
    const double* Graw = A + K * K;
    const double* dscale = Graw + K * K;
    double* invd = A + 3 * K * K + K;   // [K] reciprocals of the factor's diagonal: the substitutions multiply instead of dividing
    double* S = invd + K;
    double* col = S + K * K;
    const int n_tri = K * (K + 1) / 2;
    double a[E];
    int ik[E];   // row | (column + 1) << 8; 0 = no element (row 0, "column -1": never published, never updated)
#pragma unroll
    for (int e = 0; e < E; e++) {
        const int idx = tid + e * nthr;
        a[e] = 0.0; ik[e] = 0;
        if (idx < n_tri) {
            const int i = tri_row(idx), k = idx - i * (i + 1) / 2;
            ik[e] = i | ((k + 1) << 8);
            a[e] = Graw[i * K + k] * dscale[i] * dscale[k];
        }
    }
    for (int j = 0; j < K; j++) {
        double* cb = col + (j & 1) * K;
#pragma unroll
        for (int e = 0; e < E; e++) if ((ik[e] >> 8) == j + 1) cb[ik[e] & 255] = a[e];
        __syncthreads();   // (two buffers: the column after next is published only after everybody has passed the next barrier)
        const double piv = cb[j];
        const bool dj = dropped[j] != 0 || !(piv > sp.pivot_tol);
        const double rinv = dj ? 0.0 : __drcp_rn(piv);
        // branch free: every element loads its two column entries (a finished or absent element reads valid addresses and keeps
        // its value) — with a branch per element the seventeen updates of a thread ran one after the other, 1 us per column
        const double* cbm1 = cb - 1;
#pragma unroll
        for (int e = 0; e < E; e++) {
            const int i = ik[e] & 255, kp = ik[e] >> 8;
            const double ci = cb[i], ck = cbm1[kp];
            const double upd = fma(-(ci * rinv), ck, a[e]);
            a[e] = kp > j + 1 ? upd : a[e];
        }
    }
    // scaling: rs_j = 1 / sqrt(pivot_j) (0 for a dropped column, which becomes the unit column), l_ij = a_ij rs_j,
    // l_jj = pivot_j rs_j, 1 / l_jj = rs_j
#pragma unroll
    for (int e = 0; e < E; e++) {
        const int i = ik[e] & 255, k = (ik[e] >> 8) - 1;
        if (i == k) {
            const bool dj = dropped[i] != 0 || !(a[e] > sp.pivot_tol);
            dropped[i] = dj ? 1 : 0;
            invd[i] = dj ? 0.0 : rsqrt(a[e]);
        }
    }
    __syncthreads();
#pragma unroll
    for (int e = 0; e < E; e++) {
        const int i = ik[e] & 255, k = (ik[e] >> 8) - 1;
        if (k >= 0) {
            const double rs = invd[k];
            const double l = (i == k && rs == 0.0) ? 1.0 : a[e] * rs;
            A[i * K + k] = l;
            if (i != k) {
                const double ri = invd[i];
                S[i * K + k] = l * (ri == 0.0 ? 1.0 : ri);
                S[k * K + i] = l * (rs == 0.0 ? 1.0 : rs);
            }
        }
    }
    __syncthreads();
    for (int j = tid; j < K; j += nthr) if (invd[j] == 0.0) invd[j] = 1.0;
    __syncthreads();
}

template <int KT, class Red>
__device__ __forceinline__ void solve_cta(const SolveParams& sp, const Red& red, double* sm, int* dropped, const int col0, const int col_step) {
    if (!solve_stage(sp, red, sm, dropped, col0, col_step)) return;
    const int K = sp.K, G = sp.G, tid = threadIdx.x;
    double* A = sm;                 // [K][K] scaled Gram -> Cholesky factor (lower)
    double* Graw = A + K * K;       // [K][K] unscaled Gram
    double* dscale = Graw + K * K;  // [K]
    double* Tm = dscale + K;        // [K][K] preconditioner (when present)
    double* invd = Tm + K * K;      // [K]
    const double* T = sp.precond;
    // the thread's first right-hand side is fetched now: its latency hides behind the factorisation
    const int i_first = col0 + tid * col_step;
    double r0[KT];
#pragma unroll
    for (int k = 0; k < KT; k++) r0[k] = (k < K && i_first < G) ? red(k, i_first) : 0.0;
    solve_factor<(KT * (KT + 1) / 2 + 127) / 128>(sp, sm, dropped);   // CTAs of at least 128 threads
    // each thread: one right-hand side at a time, vectors in registers (KT = compile-time bound of K)
    for (int i = i_first; i < G; i += blockDim.x * col_step) {
        double x[KT], r[KT], y[KT], rhs[KT];
#pragma unroll
        for (int k = 0; k < KT; k++) { rhs[k] = (i == i_first) ? r0[k] : (k < K ? red(k, i) : 0.0); r[k] = rhs[k]; x[k] = 0.0; y[k] = 0.0; }
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
#pragma unroll
            for (int k = 0; k < KT; k++) {   // L y = D r
                if (k < K) {
                    double s = r[k] * dscale[k];
#pragma unroll
                    for (int m = 0; m < k; m++) s -= A[k * K + m] * y[m];
                    y[k] = dropped[k] ? 0.0 : s * invd[k];
                }
            }
#pragma unroll
            for (int k = KT - 1; k >= 0; k--) {   // L^T z = y
                if (k < K) {
                    double s = y[k];
#pragma unroll
                    for (int m = k + 1; m < KT; m++) if (m < K) s -= A[m * K + k] * y[m];
                    y[k] = dropped[k] ? 0.0 : s * invd[k];
                }
            }
#pragma unroll
            for (int k = 0; k < KT; k++) if (k < K) x[k] += y[k] * dscale[k];
            if (pass == 0) {   // residual of the unscaled system
#pragma unroll
                for (int k = 0; k < KT; k++) {
                    if (k < K) {
                        double s = rhs[k];
#pragma unroll
                        for (int m = 0; m < KT; m++) if (m < K) s = fma(-Graw[k * K + m], x[m], s);
                        r[k] = dropped[k] ? 0.0 : s;
                    }
                }
            }
        }
        double* b = sp.beta + (size_t)i * sp.ldb;
        if (T) {   // back to the monomial basis: beta = T x
#pragma unroll
            for (int k = 0; k < KT; k++) {
                if (k < K) {
                    double acc = 0.0;
#pragma unroll
                    for (int m = k; m < KT; m++) if (m < K) acc = fma(Tm[k * K + m], x[m], acc);
                    b[k] = acc;
                }
            }
            for (int k = K; k < sp.ldb; k++) b[k] = 0.0;
        } else {
#pragma unroll
            for (int k = 0; k < KT; k++) if (k < K) b[k] = x[k];
            for (int k = K; k < sp.ldb; k++) b[k] = 0.0;
        }
    }
}

// The same solve with one WARP per right-hand side (lane = row; rows lane and lane + 32 when K > 32): the substitutions
// advance one column per step with a shuffle broadcast of the finished unknown, so a solve costs about 7 K short steps
// instead of K^2 dependent operations of one thread.  CTA-wide call; warp w takes the right-hand sides
// col0 + w * col_step, col0 + (w + nwarps) * col_step, ...
template <int R, class Red>   // R = rows per lane: 1 (K <= 32) or 2 (K <= 64)
__device__ __forceinline__ void solve_cta_warps(const SolveParams& sp, const Red& red, double* sm, int* dropped, const int col0, const int col_step,
                                                long long* tr = nullptr, const bool precond_staged = false) {
    if (!solve_stage(sp, red, sm, dropped, col0, col_step, precond_staged)) return;
    if (tr) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tr[3]));
    const int K = sp.K, G = sp.G, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double* A = sm;
    const double* Graw = A + K * K;
    const double* dscale = Graw + K * K;
    const double* Tm = dscale + K;
    const double* invd = Tm + K * K;
    const double* S = invd + K;     // rows of the factor scaled by 1 / l_jj, both triangles (solve_factor)
    const bool hasT = sp.precond != nullptr;
    const int i_first = col0 + warp * col_step;
    double rhs[R];
#pragma unroll
    for (int q = 0; q < R; q++) { const int k = lane + 32 * q; rhs[q] = (k < K && i_first < G) ? red(k, i_first) : 0.0; }
    solve_factor<(32 * R * (32 * R + 1) / 2 + 127) / 128>(sp, sm, dropped);   // CTAs of at least 128 threads
    if (tr) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tr[4]));
    for (int i = i_first; i < G; i += nwarps * col_step) {
        if (i != i_first) {
#pragma unroll
            for (int q = 0; q < R; q++) { const int k = lane + 32 * q; rhs[q] = k < K ? red(k, i) : 0.0; }
        }
        double x[R], r[R], sc[R], iv[R];
        bool dr[R];
#pragma unroll
        for (int q = 0; q < R; q++) {
            const int k = lane + 32 * q;
            x[q] = 0.0; r[q] = rhs[q];
            sc[q] = k < K ? dscale[k] : 0.0; iv[q] = k < K ? invd[k] : 0.0; dr[q] = k < K ? dropped[k] != 0 : true;
        }
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
            // Both substitutions in "unit diagonal" form: the column (row) of L is scaled by 1 / l_jj when it is applied, so the value
            // that travels down the dependency chain is the lane's running right-hand side itself — one shuffle and one FMA per
            // step — and the division by the diagonal happens once per lane after the loop.
            double b[R], y[R];
#pragma unroll
            for (int q = 0; q < R; q++) b[q] = r[q] * sc[q];
            // L y = D r, column by column
#pragma unroll
            for (int qj = 0; qj < R; qj++) {
#pragma unroll 4
                for (int j = 32 * qj; j < min(K, 32 * qj + 32); j++) {
                    const double zj = __shfl_sync(0xffffffffu, dr[qj] ? 0.0 : b[qj], j & 31);
#pragma unroll
                    for (int q = 0; q < R; q++) {
                        const int k = lane + 32 * q;
                        const double lkj = (k > j && k < K) ? S[j * K + k] : 0.0;   // l_kj / l_jj
                        b[q] = fma(-lkj, zj, b[q]);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < R; q++) y[q] = dr[q] ? 0.0 : b[q] * iv[q];
            // L^T z = y, from the last column up; z overwrites y
#pragma unroll
            for (int qj = R - 1; qj >= 0; qj--) {
#pragma unroll 4
                for (int j = min(K, 32 * qj + 32) - 1; j >= 32 * qj; j--) {
                    const double wj = __shfl_sync(0xffffffffu, dr[qj] ? 0.0 : y[qj], j & 31);
#pragma unroll
                    for (int q = 0; q < R; q++) {
                        const int k = lane + 32 * q;
                        const double ljk = (k < j) ? S[j * K + k] : 0.0;   // l_jk / l_jj
                        y[q] = fma(-ljk, wj, y[q]);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < R; q++) y[q] = dr[q] ? 0.0 : y[q] * iv[q];
#pragma unroll
            for (int q = 0; q < R; q++) x[q] += y[q] * sc[q];
            if (pass == 0) {   // residual of the unscaled system
#pragma unroll
                for (int q = 0; q < R; q++) r[q] = rhs[q];
#pragma unroll
                for (int qm = 0; qm < R; qm++)
                    for (int m = 32 * qm; m < min(K, 32 * qm + 32); m++) {
                        const double xm = __shfl_sync(0xffffffffu, x[qm], m & 31);
#pragma unroll
                        for (int q = 0; q < R; q++) { const int k = lane + 32 * q; if (k < K) r[q] = fma(-Graw[k * K + m], xm, r[q]); }
                    }
#pragma unroll
                for (int q = 0; q < R; q++) if (dr[q]) r[q] = 0.0;
            }
        }
        double out[R];
#pragma unroll
        for (int q = 0; q < R; q++) out[q] = hasT ? 0.0 : x[q];
        if (hasT) {   // back to the monomial basis: beta = T x (T upper triangular)
#pragma unroll
            for (int qm = 0; qm < R; qm++)
                for (int m = 32 * qm; m < min(K, 32 * qm + 32); m++) {
                    const double xm = __shfl_sync(0xffffffffu, x[qm], m & 31);
#pragma unroll
                    for (int q = 0; q < R; q++) { const int k = lane + 32 * q; if (k <= m && k < K) out[q] = fma(Tm[k * K + m], xm, out[q]); }
                }
        }
        double* bp = sp.beta + (size_t)i * sp.ldb;
#pragma unroll
        for (int q = 0; q < R; q++) { const int k = lane + 32 * q; if (k < sp.ldb) bp[k] = k < K ? out[q] : 0.0; }
    }
}

}  // namespace lsmc
