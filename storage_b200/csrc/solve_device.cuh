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

inline size_t solve_smem_bytes(int K) { return sizeof(double) * ((size_t)3 * K * K + 2 * K); }

// Gram matrix -> shared memory, equilibration, preconditioner.  Returns false in the current-period mode (means, no system).
template <class Red>
__device__ __forceinline__ bool solve_stage(const SolveParams& sp, const Red& red, double* sm, int* dropped, const int col0, const int col_step) {
    const int K = sp.K, G = sp.G, tid = threadIdx.x;
    if (sp.accum_mode == 2) {   // current period: continuation = mean over sims (LSV.cs:171-182)
        for (int i = col0 + tid * col_step; i < G; i += blockDim.x * col_step) {
            double* b = sp.beta + (size_t)i * sp.ldb;
            b[0] = red(0, i) / (double)sp.n_global;
            for (int k = 1; k < sp.ldb; k++) b[k] = 0.0;
        }
        return false;
    }
    double* A = sm;                 // [K][K] scaled Gram -> Cholesky factor (lower)
    double* Graw = A + K * K;       // [K][K] unscaled Gram
    double* dscale = Graw + K * K;  // [K]
    double* Tm = dscale + K;        // [K][K] preconditioner (when present)
    const int gcol = 8 * sp.NT;
    const double* T = sp.precond;
#pragma unroll 4   // independent L2 loads: several in flight per thread (K = 56: 25 elements per thread, 22 us -> a few)
    for (int e = tid; e < K * K; e += blockDim.x) {
        int r = e / K, c = e % K;
        // symmetrise: the two triangles come from different DMMA tiles and may differ in the last bit
        Graw[e] = 0.5 * (red(r, gcol + c) + red(c, gcol + r));
    }
    if (T) {
#pragma unroll 4
        for (int e = tid; e < K * K; e += blockDim.x) Tm[e] = T[e];
    }
    __syncthreads();
    for (int k = tid; k < K; k += blockDim.x) {
        const double gkk = Graw[k * K + k];
        const bool flagged = T && T[(size_t)K * K + k] != 0.0;   // column found dependent by the preconditioner's QR
        dscale[k] = (gkk > 0.0 && !flagged) ? rsqrt(gkk) : 0.0;
        dropped[k] = (gkk > 0.0 && !flagged) ? 0 : 1;
    }
    __syncthreads();
    for (int e = tid; e < K * K; e += blockDim.x) {
        int r = e / K, c = e % K;
        A[e] = Graw[e] * dscale[r] * dscale[c];
    }
    __syncthreads();
    return true;
}

// Cholesky of the staged, scaled Gram matrix by the whole CTA + reciprocals of the diagonal (CTA-wide call).
// Right-looking with the scaling deferred: at column j every thread reads the pivot itself and applies
//     A[i][k] -= A[i][j] A[k][j] / pivot          (j < k <= i)
// to its share of the trailing triangle — the reads are in column j, the writes in the columns right of it, so one barrier
// per column suffices and nobody waits for a thread that takes a square root; L = A scaled by 1 / sqrt(pivot) comes in one
// pass at the end.  (The previous form — warp 0 alone, sqrt and divisions serial on lane 0, three warp barriers per column —
// took 5.8 us at K = 10 and 104 us at K = 56: tail trace in profiles/.)
__device__ __forceinline__ void solve_factor(const SolveParams& sp, double* sm, int* dropped) {
    const int K = sp.K, tid = threadIdx.x, nthr = blockDim.x;
    double* A = sm;
    double* Tm = A + 2 * K * K + K;
    double* invd = Tm + K * K;      // [K] reciprocals of the factor's diagonal: the substitutions multiply instead of dividing
    const int ti = tid >> 4, tk = tid & 15, ni = nthr >> 4;   // blockDim.x is a multiple of 32
    for (int j = 0; j < K; j++) {
        const double piv = A[j * K + j];
        const bool dj = dropped[j] != 0 || !(piv > sp.pivot_tol);
        const double rinv = dj ? 0.0 : __drcp_rn(piv);
        for (int i = j + 1 + ti; i < K; i += ni) {
            const double aij = A[i * K + j] * rinv;
            for (int k = j + 1 + tk; k <= i; k += 16) A[i * K + k] = fma(-aij, A[k * K + j], A[i * K + k]);
        }
        __syncthreads();
    }
    // scaling in two passes: rs_j = 1 / sqrt(pivot_j) (0 for a dropped column, which becomes the unit column), then
    // l_ij = a_ij rs_j, l_jj = pivot_j rs_j, 1 / l_jj = rs_j
    for (int j = tid; j < K; j += nthr) {
        const double piv = A[j * K + j];
        const bool dj = dropped[j] != 0 || !(piv > sp.pivot_tol);
        dropped[j] = dj ? 1 : 0;
        invd[j] = dj ? 0.0 : rsqrt(piv);
    }
    __syncthreads();
    for (int e = tid; e < K * K; e += nthr) {
        const int i = e / K, j = e - i * K;
        if (j > i) continue;
        const double rs = invd[j];
        A[e] = (i == j) ? (rs == 0.0 ? 1.0 : A[e] * rs) : A[e] * rs;
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
    solve_factor(sp, sm, dropped);
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
                                                long long* tr = nullptr) {
    if (!solve_stage(sp, red, sm, dropped, col0, col_step)) return;
    if (tr) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tr[3]));
    const int K = sp.K, G = sp.G, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double* A = sm;
    const double* Graw = A + K * K;
    const double* dscale = Graw + K * K;
    const double* Tm = dscale + K;
    const double* invd = Tm + K * K;
    const bool hasT = sp.precond != nullptr;
    const int i_first = col0 + warp * col_step;
    double rhs[R];
#pragma unroll
    for (int q = 0; q < R; q++) { const int k = lane + 32 * q; rhs[q] = (k < K && i_first < G) ? red(k, i_first) : 0.0; }
    solve_factor(sp, sm, dropped);
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
                    const double ivj = invd[j];
#pragma unroll
                    for (int q = 0; q < R; q++) {
                        const int k = lane + 32 * q;
                        const double lkj = (k > j && k < K) ? A[k * K + j] * ivj : 0.0;
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
                    const double ivj = invd[j];
#pragma unroll
                    for (int q = 0; q < R; q++) {
                        const int k = lane + 32 * q;
                        const double ljk = (k < j) ? A[j * K + k] * ivj : 0.0;
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
