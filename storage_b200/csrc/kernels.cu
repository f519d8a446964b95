// kernels.cu — hand-written sm_100a kernels of the LSMC hot path.
//   K1 simulate_kernel       multi-factor OU simulation (LVP.cs:160-178 -> Cmdty.Core simulator)
//   K3 sweep_kernel          backward decision sweep of one period fused with the design-matrix build and the
//                            X^T X / X^T Y normal-equation accumulation of the previous period on FP64 tensor
//                            cores (DMMA) (LSV.cs:185-205, 234-332, 753-770)
//   K2 reduce/solve kernels  partial reduction + scaled Cholesky solve of the K x K system for all grid points
//   K4 forward_kernel        forward inventory / cash-flow simulation of one period (LSV.cs:426-480)
#include "kernels.h"
#include "solve_device.cuh"

#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

namespace lsmc {

cudaError_t ensure_dynamic_smem(const void* kernel, size_t bytes) {
    static std::mutex mu;   // (no shortcut below 48 KB: the default limit counts the kernel's static shared memory too)
    static std::map<std::pair<const void*, int>, size_t> granted;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    size_t& have = granted[std::make_pair(kernel, dev)];
    if (bytes <= have) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) have = bytes;
    return e;
}

bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("LSMC_PDL"); return !(e && e[0] == '0'); }();
    return on;
}

// =================================================================================================
// K1: simulation.  One thread per path, the time loop runs in registers; stores are coalesced across the warp
// (path-major rows).  HBM-bound on 8(F+1) bytes per path-step written (+8F read with injected normals).
// =================================================================================================
template <int F>
__global__ void __launch_bounds__(256) simulate_kernel(SimParams sp) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= sp.N) return;
    double x[F];
#pragma unroll
    for (int f = 0; f < F; f++) x[f] = 0.0;
    const int64_t gn = sp.sim_offset + n;
    double sign = 1.0;
    uint64_t draw = (uint64_t)gn;
    if (sp.mode == LSMC_RNG_PHILOX && sp.antithetic) { draw = (uint64_t)(gn >> 1); sign = (gn & 1) ? -1.0 : 1.0; }   // partners are adjacent paths
    for (int j = 0; j < sp.S; j++) {
        double z[F + 1];
        if (sp.mode == LSMC_RNG_NORMALS) {
#pragma unroll
            for (int f = 0; f < F; f++) z[f] = __ldcs(sp.normals + ((size_t)j * F + f) * sp.N + n);
        } else {
#pragma unroll
            for (int f = 0; f < F; f += 2) {
                double a, b;
                philox_normal_pair(sp.seed, draw, (uint32_t)j, (uint32_t)(f / 2), a, b);
                z[f] = sign * a;
                z[f + 1] = sign * b;
            }
        }
        const double* L = sp.chol + (size_t)j * F * F;
        double acc = 0.0;
#pragma unroll
        for (int f = 0; f < F; f++) {
            double eps = mul_rn(__ldg(L + f * F), z[0]);
#pragma unroll
            for (int b = 1; b <= f; b++) eps = add_rn(eps, mul_rn(__ldg(L + f * F + b), z[b]));
            x[f] = add_rn(mul_rn(__ldg(sp.decay + (size_t)j * F + f), x[f]), eps);
        }
#pragma unroll
        for (int f = 0; f < F; f++) {
            double term = mul_rn(__ldg(sp.vol + (size_t)j * F + f), x[f]);
            acc = (f == 0) ? term : add_rn(acc, term);
            __stcs(sp.fac + ((size_t)j * F + f) * sp.Npad + n, x[f]);
        }
        double s = mul_rn(__ldg(sp.fwd + j), det_exp(sub_rn(acc, __ldg(sp.drift + j))));
        __stcs(sp.spot + (size_t)j * sp.Npad + n, s);
    }
}

// Antithetic device-RNG form: one thread per pair of partner paths (2m, 2m+1).  The partner's draws are the exact
// negatives, so its factors are the exact negatives too (every operation of the OU update is odd in the draws): one
// Philox + Box-Muller evaluation and one state vector serve both paths, only the exponential is evaluated twice.
// Bit-identical to simulate_kernel on the same seed.
template <int F>
__global__ void __launch_bounds__(256) simulate_pair_kernel(SimParams sp) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = 2 * m;
    if (n >= sp.N) return;
    double x[F];
#pragma unroll
    for (int f = 0; f < F; f++) x[f] = 0.0;
    const uint64_t draw = (uint64_t)((sp.sim_offset + n) >> 1);
    for (int j = 0; j < sp.S; j++) {
        double z[F + 1];
#pragma unroll
        for (int f = 0; f < F; f += 2) philox_normal_pair(sp.seed, draw, (uint32_t)j, (uint32_t)(f / 2), z[f], z[f + 1]);
        const double* L = sp.chol + (size_t)j * F * F;
        double acc = 0.0;
#pragma unroll
        for (int f = 0; f < F; f++) {
            double eps = mul_rn(__ldg(L + f * F), z[0]);
#pragma unroll
            for (int b = 1; b <= f; b++) eps = add_rn(eps, mul_rn(__ldg(L + f * F + b), z[b]));
            x[f] = add_rn(mul_rn(__ldg(sp.decay + (size_t)j * F + f), x[f]), eps);
        }
#pragma unroll
        for (int f = 0; f < F; f++) {
            double term = mul_rn(__ldg(sp.vol + (size_t)j * F + f), x[f]);
            acc = (f == 0) ? term : add_rn(acc, term);
            __stcs(reinterpret_cast<double2*>(sp.fac + ((size_t)j * F + f) * sp.Npad + n), make_double2(x[f], -x[f]));
        }
        const double drift = __ldg(sp.drift + j), fwd = __ldg(sp.fwd + j);
        const double s0 = mul_rn(fwd, det_exp(sub_rn(acc, drift)));
        const double s1 = mul_rn(fwd, det_exp(sub_rn(-acc, drift)));
        __stcs(reinterpret_cast<double2*>(sp.spot + (size_t)j * sp.Npad + n), make_double2(s0, s1));
    }
}

void launch_simulate(const SimParams& sp, cudaStream_t stream) {
    const int threads = 256;
    const bool pairs = sp.mode == LSMC_RNG_PHILOX && sp.antithetic && (sp.N % 2 == 0) && (sp.sim_offset % 2 == 0) && (sp.Npad % 2 == 0);
    if (pairs) {
        const int blocks = (int)((sp.N / 2 + threads - 1) / threads);
        switch (sp.F) {
#define CASE(f) case f: simulate_pair_kernel<f><<<blocks, threads, 0, stream>>>(sp); break;
            CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
        }
        return;
    }
    const int blocks = (int)((sp.N + threads - 1) / threads);
    switch (sp.F) {
#define CASE(f) case f: simulate_kernel<f><<<blocks, threads, 0, stream>>>(sp); break;
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    }
}

// K3 (the fused backward sweep) lives in sweep.cu.

// =================================================================================================
// K2p: regression preconditioner.  One CTA per simulated period: Householder QR (no pivoting, like the reference's
// thin QR, LSV.cs:187) of the design matrix on the first `nsub` paths, held column-major in shared memory; writes
// T = R^-1 and the dropped-column flags.  The sweep accumulates the normal equations in the basis x~ = T^T x, whose
// Gram matrix is ~ (N/nsub) I, so the Cholesky solve that follows has QR-grade accuracy however ill-conditioned the
// raw monomials are (cond(X) reaches 1e13 at the first half-hourly steps of a two-factor model).
// =================================================================================================
__device__ __forceinline__ double block_sum(double v, double* scratch) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    double tot = 0.0;
    for (int w = 0; w < nw; w++) tot += scratch[w];
    return tot;
}

__global__ void __launch_bounds__(512) precond_kernel(PrecondParams q) {
    extern __shared__ double psm[];
    const int K = q.K, n = q.nsub, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    const int j = blockIdx.x;
    const BasisDev* __restrict__ bd = q.basis;
    double* A = psm;                      // [K][n] column-major
    double* R = A + (size_t)K * n;        // [K][K]
    double* col0 = R + K * K;             // [K] original column norms
    double* scratch = col0 + K;           // [32]
    __shared__ int s_dropped[64];
    __shared__ double s_tau, s_scale;
    __shared__ int s_skip;

    for (int i = tid; i < n; i += blockDim.x) {
        const double s = q.spot[(size_t)j * q.Npad + i];
        double f[kMaxFactors];
        for (int a = 0; a < q.F; a++) f[a] = q.fac[((size_t)j * q.F + a) * q.Npad + i];
        for (int k = 0; k < K; k++) {
            double acc = 1.0;
            for (int e = 0; e < bd->spot_pow[k]; e++) acc *= s;
            for (int a = 0; a < q.F; a++)
                for (int e = 0; e < bd->fac_pow[k][a]; e++) acc *= f[a];
            A[(size_t)k * n + i] = acc;
        }
    }
    for (int e = tid; e < K * K; e += blockDim.x) R[e] = 0.0;
    __syncthreads();
    for (int k = 0; k < K; k++) {
        double part = 0.0;
        for (int i = tid; i < n; i += blockDim.x) { const double a = A[(size_t)k * n + i]; part += a * a; }
        const double tot = block_sum(part, scratch);
        if (tid == 0) col0[k] = sqrt(tot);
    }
    __syncthreads();

    int r = 0;   // pivot row = number of columns kept so far (uniform across the block)
    for (int k = 0; k < K; k++) {
        double* ck = A + (size_t)k * n;
        double part = 0.0;
        for (int i = r + tid; i < n; i += blockDim.x) part += ck[i] * ck[i];
        const double normsq = block_sum(part, scratch);
        if (tid == 0) {
            const double norm = sqrt(normsq);
            if (r >= n || !(norm > 1e-13 * col0[k]) || col0[k] == 0.0) {
                s_skip = 1; s_dropped[k] = 1;
            } else {
                const double alpha = ck[r];
                const double beta = alpha >= 0.0 ? -norm : norm;
                s_tau = (beta - alpha) / beta;
                s_scale = 1.0 / (alpha - beta);
                s_skip = 0; s_dropped[k] = 0;
                R[k * K + k] = beta;
            }
        }
        __syncthreads();
        if (s_skip) {
            if (tid == 0) { for (int m = 0; m < K; m++) { R[m * K + k] = 0.0; } R[k * K + k] = 1.0; }
            __syncthreads();
            continue;
        }
        const double tau = s_tau, scale = s_scale;
        for (int i = r + 1 + tid; i < n; i += blockDim.x) ck[i] *= scale;
        __syncthreads();
        // apply H = I - tau v v^T (v_r = 1) to the trailing columns: one warp per column
        for (int c = k + 1 + warp; c < K; c += nw) {
            double* cc = A + (size_t)c * n;
            double dot = 0.0;
            for (int i = r + 1 + lane; i < n; i += 32) dot += ck[i] * cc[i];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, off);
            dot = (dot + cc[r]) * tau;
            for (int i = r + 1 + lane; i < n; i += 32) cc[i] -= dot * ck[i];
            if (lane == 0) { cc[r] -= dot; R[k * K + c] = cc[r]; }
        }
        __syncthreads();
        r++;
    }
    // rows of R that belong to dropped columns stay e_k; entries R[k][c] for a dropped c were zeroed when c was
    // skipped only for m < c processed so far -> clear the whole column again now
    for (int e = tid; e < K * K; e += blockDim.x) {
        const int row = e / K, c = e % K;
        if (row != c && (s_dropped[c] || s_dropped[row])) R[e] = 0.0;
    }
    __syncthreads();
    // T = R^-1, one thread per column
    double* out = q.out + (size_t)j * precond_stride(K);
    if (tid < K) {
        const int c = tid;
        double tcol[64];
        for (int i = K - 1; i >= 0; i--) {
            if (i > c) { tcol[i] = 0.0; continue; }
            double sum = (i == c) ? 1.0 : 0.0;
            for (int m = i + 1; m <= c; m++) sum -= R[i * K + m] * tcol[m];
            tcol[i] = sum / R[i * K + i];
        }
        for (int i = 0; i < K; i++) out[(size_t)i * K + c] = tcol[i];
        out[(size_t)K * K + c] = s_dropped[c] ? 1.0 : 0.0;
    }
}

int precond_rows(int K, int64_t N) {
    int cap = (int)((200 * 1024 / sizeof(double) - (size_t)K * K - K - 32) / K);
    cap = std::min(cap, 2048) & ~31;
    return (int)std::min<int64_t>(cap, N);
}

void launch_precond(const PrecondParams& q, cudaStream_t stream) {
    const size_t smem = sizeof(double) * ((size_t)q.K * q.nsub + (size_t)q.K * q.K + q.K + 32);
    if (ensure_dynamic_smem((const void*)precond_kernel, smem) != cudaSuccess) return;   // the launch error surfaces through cudaGetLastError
    precond_kernel<<<q.S, 512, smem, stream>>>(q);
}

// =================================================================================================
// K2a: deterministic reduction of the per-CTA partials: one warp per element, fixed lane/CTA order.
// =================================================================================================
__global__ void __launch_bounds__(256) reduce_partials_kernel(const double* __restrict__ partial, int n_parts, int n_elem,
                                                              double* __restrict__ out) {
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (e >= n_elem) return;
    double s = 0.0;
    for (int c = lane; c < n_parts; c += 32) s += partial[(size_t)c * n_elem + e];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) out[e] = s;
}

void launch_reduce_partials(const double* partial, int n_parts, int n_elem, double* out, cudaStream_t stream) {
    const int wpb = 8;
    reduce_partials_kernel<<<(n_elem + wpb - 1) / wpb, wpb * 32, 0, stream>>>(partial, n_parts, n_elem, out);
}

// =================================================================================================
// K2b: solve the normal equations for all grid points of one period.
//   reduced[k][col]: col < G -> X^T V_i ; col = 8*NT + k' -> X^T X.
// Column-equilibrated Cholesky (G~ = D G D, D = diag(G)^-1/2) with a pivot guard that drops numerically dependent
// columns, one step of iterative refinement against the unscaled system, beta = D x.
// =================================================================================================
template <int KT>
__global__ void __launch_bounds__(128) solve_kernel(SolveParams sp) {
    extern __shared__ double sm[];
    __shared__ int dropped[64];
    solve_cta<KT>(sp, GlobalRed{sp.reduced, sp.ncols}, sm, dropped, 0, 1);
}

void launch_solve(const SolveParams& sp, cudaStream_t stream) {
    size_t smem = solve_smem_bytes(sp.K);
    if (sp.K <= 4) solve_kernel<4><<<1, 128, smem, stream>>>(sp);
    else if (sp.K <= 8) solve_kernel<8><<<1, 128, smem, stream>>>(sp);
    else if (sp.K <= 12) solve_kernel<12><<<1, 128, smem, stream>>>(sp);
    else if (sp.K <= 16) solve_kernel<16><<<1, 128, smem, stream>>>(sp);
    else {
#ifndef LSMC_FAST_BUILD
        if (ensure_dynamic_smem((const void*)solve_kernel<64>, smem) != cudaSuccess) return;
        solve_kernel<64><<<1, 128, smem, stream>>>(sp);
#endif
    }
}

// =================================================================================================
// K4: forward simulation of one period (LSV.cs:426-480).  One thread per path (grid-stride), state = inventory and
// accumulated PV; only the <= 2 fitted values per decision that are actually needed are evaluated (x.beta_l on
// demand) instead of the reference's full N x G product (LSV.cs:395-401).  Block-level sums feed the expected
// profile, deltas and trigger prices.
// =================================================================================================
__device__ __forceinline__ int last_at_or_below(const double* x, int n, double v) {
    int lo = 0, hi = n;   // first index with x > v
    while (lo < hi) { int mid = (lo + hi) >> 1; if (__ldg(x + mid) <= v) lo = mid + 1; else hi = mid; }
    return lo - 1;
}

// Polynomial.Evaluate (PolynomialInjectWithdrawConstraint.cs:80-85): Horner from the highest coefficient, no contraction.
__device__ __forceinline__ double horner_dev(const double* coef, int n, double z) {
    double sum = __ldg(coef + n - 1);
    for (int i = n - 2; i >= 0; i--) sum = add_rn(mul_rn(sum, z), __ldg(coef + i));
    return sum;
}

__device__ __forceinline__ double spline_dev(const double* x, const double* y, int n, double v) {
    int k = min(max(last_at_or_below(x, n, v), 0), n - 2);
    double x0 = __ldg(x + k), y0 = __ldg(y + k);
    double slope = (__ldg(y + k + 1) - y0) / (__ldg(x + k + 1) - x0);
    return add_rn(y0, mul_rn(sub_rn(v, x0), slope));
}

#ifndef LSMC_FWD_MINB
#define LSMC_FWD_MINB 3
#endif
constexpr int kFwdThreads = 256;

// All periods [p0, p1) in one launch: a thread owns its paths for the whole pass (no cross-path dependence in the forward
// simulation, LSV.cs:426-480), the per-period descriptors come from a device array.  Per period the CTA stages the
// coefficient rows beta[Gn][K] (odd pitch: the rows a warp's lanes need are scattered over the grid) and the next grid in
// shared memory; basis-program nodes live in a per-thread column of shared memory, the regressor row and the column sums
// in registers (KT = compile-time bound of K; EXACT: K == KT).
// StorageHelper.BisectInventorySpace (SH.cs:280-314) on the staged grid; returns false when outside the grid.
__device__ __forceinline__ bool bisect_sm(const double* grid, int n, double inv, double tol, int& lo, int& hi) {
    if (n == 1 && fabs(inv - grid[0]) <= tol) { lo = hi = 0; return true; }
    int left = 0, right = n - 1;
    const int top = right;
    while (right > left) {
        int mid = (left + right) >> 1;
        double gm = grid[mid];
        if (fabs(inv - gm) <= tol) { lo = hi = mid; return true; }
        if (gm > inv) { right = mid; continue; }
        double gn = grid[mid + 1];
        if (inv <= gn) { lo = mid; hi = mid + 1; return true; }
        if (fabs(inv - gn) <= tol) { lo = hi = mid + 1; return true; }
        if (mid + 1 == top) return false;
        left = mid;
    }
    return false;
}

// The same result as bisect_sm in O(1) on a grid whose intervals are all wider than 2 tol: find the interval with
// grid[m] < inv <= grid[m+1] from the fixed spacing (guess + fix-up), then apply the bisection's tolerance rules.  The one
// path-dependent case of SH.cs:280-314 — inv within tol at or below grid[i] gives (i, i) when the bisection probes i before
// i-1 and (i-1, i) otherwise — is a function of (n, i) alone, tabulated per period in probe_first[] (fwd_probe_first).
__device__ __forceinline__ bool locate_sm(const double* grid, const unsigned char* probe_first, int n, double g0, double inv_h,
                                          double inv, double tol, int& lo, int& hi) {
    const int top = n - 1;
    if (n == 1) { lo = hi = 0; return fabs(inv - grid[0]) <= tol; }
    int m = (int)((inv - g0) * inv_h);
    m = min(max(m, 0), top - 1);
    while (m > 0 && grid[m] >= inv) m--;
    while (m < top - 1 && grid[m + 1] < inv) m++;
    const double gl = grid[m], gu = grid[m + 1];
    if (inv <= gl) { lo = hi = 0; return fabs(inv - gl) <= tol; }          // m == 0: at or below the first point
    if (inv > gu) { lo = hi = top; return fabs(inv - gu) <= tol; }         // m == top-1: above the last point
    if (fabs(inv - gu) <= tol) { hi = m + 1; lo = probe_first[m + 1] ? m + 1 : m; return true; }
    if (fabs(inv - gl) <= tol) { lo = hi = m; return true; }
    lo = m; hi = m + 1;
    return true;
}

// probe_first[i]: does the bisection of SH.cs:280-314, searching a value just below grid[i], probe index i before i-1?
__device__ __forceinline__ unsigned char fwd_probe_first(int n, int i) {
    int left = 0, right = n - 1;
    while (right > left) {
        const int mid = (left + right) >> 1;
        if (mid == i) return 1;
        if (mid > i) right = mid;
        else { if (mid == i - 1) return 0; left = mid; }
    }
    return 0;
}

// Result of the bisection of SH.cs:280-314 for every target of one comparison class.  A target t is classified by the grid
// interval that holds it — m with grid[m] < t <= grid[m+1]; m = -1: t <= grid[0]; m = n-1: t > grid[n-1] — and by which of
// the points grid[m], grid[m+1], grid[n-2], grid[n-1] it is within the tolerance of (bits 0..3 of flags).  On grids whose
// intervals are all wider than twice the tolerance, except possibly the last one (FixedSpacingStateSpaceGridCalc.cs:45-62
// accumulates steps and then clamps to the maximum, which can leave a last interval of one rounding error), no other point
// can be within the tolerance, so every comparison the bisection makes is determined by the class and its control flow can
// be replayed without the target: lo | hi << 8, or 0xffff where it reports "outside the grid".
__device__ __forceinline__ uint16_t fwd_class_result(int n, int m, int flags) {
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

__host__ __device__ constexpr int fwd_beta_pitch(int KT) { return KT | 1; }
constexpr int kFwdPillars = 16;
constexpr int kFwdMaxGrid = 127;   // forward_fast_kernel: class table of (grid points + 1) x 16 entries   // piecewise-linear ratchet tables up to this size are staged in shared memory

template <int KT, bool EXACT>
__global__ void __launch_bounds__(kFwdThreads, KT <= 16 ? LSMC_FWD_MINB : 1) forward_kernel(const ForwardParams* __restrict__ pp, int p0, int p1, int max_grid) {
    extern __shared__ double fsm[];   // [n_nodes][blockDim] node columns, [max_grid][BP] beta, [max_grid] grid, [warps][nsum] reduction scratch
    __shared__ ForwardParams sq;
    __shared__ uint32_t s_ops[kMaxOps];              // basis program: dst | a << 8 | b << 16
    __shared__ uint8_t s_node_of_basis[64];
    __shared__ unsigned char s_probe_first[256];
    __shared__ double s_cx[kFwdPillars], s_clo[kFwdPillars], s_chi[kFwdPillars], s_slo[kFwdPillars], s_shi[kFwdPillars];
    constexpr int BP = fwd_beta_pitch(KT);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    {
        const BasisDev* __restrict__ bd0 = pp[p0].basis;
        for (int o = tid; o < bd0->n_ops; o += blockDim.x) {
            const BasisOpDev op = bd0->ops[o];
            s_ops[o] = (uint32_t)op.dst | ((uint32_t)op.a << 8) | ((uint32_t)op.b << 16);
        }
        for (int k = tid; k < 64; k += blockDim.x) s_node_of_basis[k] = (uint8_t)bd0->node_of_basis[k];
    }
    for (int p = p0; p < p1; p++) {
        __syncthreads();
        if (tid == 0) sq = pp[p];
        __syncthreads();
        const ForwardParams& q = sq;
        const BasisDev* __restrict__ bd = q.basis;
        const int K = EXACT ? KT : q.K, nsum = kFwdSums + K;
        const int n_ops = bd->n_ops;
        double* nodes = fsm + tid;                                   // node j at nodes[j * blockDim.x]
        double* beta_s = fsm + (size_t)bd->n_nodes * blockDim.x;     // row l at beta_s[l * BP]
        double* grid_s = beta_s + (size_t)max_grid * BP;
        double* red = grid_s + max_grid;                             // [warps][nsum]
        const int Gn = q.Gn;
        int narrow = 0;   // an interval of the next grid within 2 tol (or more rows than the table holds): plain bisection
        {
            const int nk = q.cont_mode == 1 ? 1 : K;
            for (int l = warp; l < Gn; l += nwarps)
                for (int k = lane; k < nk; k += 32) beta_s[l * BP + k] = __ldg(q.beta + (size_t)l * q.ldb + k);
            for (int l = tid; l < Gn; l += blockDim.x) {
                grid_s[l] = __ldg(q.next_grid + l);
                if (l < 256) s_probe_first[l] = fwd_probe_first(Gn, l);
                if (l + 1 < Gn && !(__ldg(q.next_grid + l + 1) - __ldg(q.next_grid + l) > 2.0 * q.tol)) narrow = 1;
            }
            if (Gn > 256) narrow = 1;
            if (q.c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR && q.c_n <= kFwdPillars && tid < q.c_n) {
                const double x0 = __ldg(q.c_x + tid), l0 = __ldg(q.c_lo + tid), h0 = __ldg(q.c_hi + tid);
                s_cx[tid] = x0; s_clo[tid] = l0; s_chi[tid] = h0;
                if (tid + 1 < q.c_n) {   // slopes as spline_dev forms them
                    const double dx = __ldg(q.c_x + tid + 1) - x0;
                    s_slo[tid] = (__ldg(q.c_lo + tid + 1) - l0) / dx;
                    s_shi[tid] = (__ldg(q.c_hi + tid + 1) - h0) / dx;
                }
            }
        }
        const bool fast_locate = __syncthreads_or(narrow) == 0;
        const bool staged_pillars = q.c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR && q.c_n <= kFwdPillars && q.c_n >= 2;
        const double g0 = grid_s[0], inv_h = Gn > 1 ? 1.0 / (grid_s[1] - grid_s[0]) : 0.0;
        double sums[kFwdSums], xsum[KT];
#pragma unroll
        for (int i = 0; i < kFwdSums; i++) sums[i] = 0.0;
#pragma unroll
        for (int k = 0; k < KT; k++) xsum[k] = 0.0;

        const int64_t stride = (int64_t)gridDim.x * blockDim.x;
        for (int64_t n = (int64_t)blockIdx.x * blockDim.x + tid; n < q.N; n += stride) {
            const double s = q.spot_is_const ? q.spot_const : __ldcs(q.spot + n);
            const double v = q.inventory[n];
            double xrow[KT];
            if (q.cont_mode == 0) {
                nodes[0] = 0.0; nodes[(size_t)blockDim.x] = 1.0; nodes[(size_t)2 * blockDim.x] = s;
                for (int f = 0; f < q.F; f++) nodes[(size_t)(3 + f) * blockDim.x] = __ldcs(q.fac + (size_t)f * q.Npad + n);
                for (int o = 0; o < n_ops; o++) {
                    const uint32_t op = s_ops[o];
                    nodes[(size_t)(op & 0xffu) * blockDim.x] = mul_rn(nodes[(size_t)((op >> 8) & 0xffu) * blockDim.x], nodes[(size_t)(op >> 16) * blockDim.x]);
                }
#pragma unroll
                for (int k = 0; k < KT; k++) {
                    xrow[k] = (k < K) ? nodes[(size_t)s_node_of_basis[k] * blockDim.x] : 0.0;
                    if (k < K) xsum[k] += xrow[k];
                }
            }
            // CmdtyStorage.GetInjectWithdrawRange (CmdtyStorage.cs:86-100)
            if (v < q.min_inv - q.slack_lo || v > q.max_inv + q.slack_hi) atomicExch(q.error_flag, 1);
            double rmin, rmax;
            if (q.c_kind == LSMC_CONSTRAINT_CONSTANT) { rmin = __ldg(q.c_lo); rmax = __ldg(q.c_hi); }
            else if (staged_pillars) {   // spline_dev on the staged table: segment by counting, slopes precomputed
                int k = 0;
                for (int i = 1; i < q.c_n - 1; i++) k += (s_cx[i] <= v) ? 1 : 0;
                const double dxv = sub_rn(v, s_cx[k]);
                rmax = add_rn(s_chi[k], mul_rn(dxv, s_shi[k]));
                rmin = add_rn(s_clo[k], mul_rn(dxv, s_slo[k]));
            }
            else if (q.c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) { rmax = spline_dev(q.c_x, q.c_hi, q.c_n, v); rmin = spline_dev(q.c_x, q.c_lo, q.c_n, v); }
            else if (q.c_kind == LSMC_CONSTRAINT_POLYNOMIAL) { rmax = horner_dev(q.c_hi, q.c_n, v); rmin = horner_dev(q.c_lo, q.c_n, v); }
            else {
                if (v < __ldg(q.c_x) - q.slack_hi || v > __ldg(q.c_x + q.c_n - 1) + q.slack_hi) atomicExch(q.error_flag, 2);
                int idx = min(max(last_at_or_below(q.c_x, q.c_n, v), 0), q.c_n - 1);
                rmin = __ldg(q.c_lo + idx); rmax = __ldg(q.c_hi + idx);
            }
            const double loss = mul_rn(q.loss_pct, v);
            // bang-bang decision set (SH.cs:109-197)
            // volume of decision d computed on demand (no per-thread array)
            int nd;
            double w, jn, stepw = 0.0, stepj = 0.0;
            bool one_sided;
            {
                const double base = sub_rn(v, loss);
                const double afterW = add_rn(rmin, base), afterJ = add_rn(rmax, base);
                if (afterW > q.next_max) { if (!(sub_rn(afterW, q.next_max) < q.tol)) atomicExch(q.error_flag, 3); w = sub_rn(q.next_max, base); }
                else w = afterW > q.next_min ? rmin : sub_rn(q.next_min, base);
                if (afterJ < q.next_min) { if (!(sub_rn(q.next_min, afterJ) < q.tol)) atomicExch(q.error_flag, 3); jn = sub_rn(q.next_min, base); }
                else jn = afterJ < q.next_max ? rmax : sub_rn(q.next_max, base);
                const int ex = q.extra;
                one_sided = (w >= 0.0 || jn <= 0.0);
                if (ex == 0) nd = one_sided ? 2 : 3;
                else if (one_sided) { nd = ex + 2; stepw = (jn - w) / (ex + 1); }
                else { nd = 2 * ex + 3; stepw = (0.0 - w) / (ex + 1); stepj = (jn - 0.0) / (ex + 1); }
            }
            auto volume_of = [&](int d) -> double {
                const int ex = q.extra;
                if (d == 0) return w;
                if (d == nd - 1) return jn;
                if (one_sided) return add_rn(mul_rn((double)d, stepw), w);
                if (d <= ex) return add_rn(mul_rn((double)d, stepw), w);
                if (d == ex + 1) return 0.0;
                return add_rn(mul_rn((double)(d - ex - 1), stepj), 0.0);
            };
            const double inv_cost_npv = q.inv_cost == 0.0 ? 0.0 : mul_rn(mul_rn(v, q.inv_cost), q.inv_cost_df);
            double best = 0.0, best_imm = 0.0, best_cons = 0.0, best_vol = 0.0;
            int best_d = 0;
            auto consider = [&](int d, double vol) {
                const double after = sub_rn(add_rn(v, vol), loss);
                const double cons = vol > 0.0 ? mul_rn(q.cons_inj, fabs(vol)) : mul_rn(q.cons_wd, fabs(-vol));
                const double a = mul_rn(mul_rn(-vol, s), q.df);
                const double c = mul_rn(mul_rn(-cons, s), q.df);
                const double cost = vol > 0.0 ? mul_rn(mul_rn(q.inj_cost, vol), q.inj_df) : mul_rn(mul_rn(q.wd_cost, fabs(-vol)), q.wd_df);
                const double imm = sub_rn(add_rn(sub_rn(a, cost), c), inv_cost_npv);
                int lo, hi;
                const bool found = fast_locate ? locate_sm(grid_s, s_probe_first, Gn, g0, inv_h, after, q.tol, lo, hi)
                                               : bisect_sm(grid_s, Gn, after, q.tol, lo, hi);
                if (!found) { atomicExch(q.error_flag, 4); lo = hi = 0; }
                auto fitted = [&](int l) {
                    const double* b = beta_s + l * BP;
                    if (q.cont_mode == 1) return b[0];
                    double acc = 0.0;
#pragma unroll
                    for (int k = 0; k < KT; k++) if (k < K) acc = fma(xrow[k], b[k], acc);
                    return acc;
                };
                double cont;
                if (lo == hi) cont = fitted(lo);
                else {
                    const double gl = grid_s[lo], gu = grid_s[hi];
                    const double wl = sub_rn(gu, after) / sub_rn(gu, gl);
                    const double wu = sub_rn(1.0, wl);
                    cont = add_rn(mul_rn(fitted(lo), wl), mul_rn(fitted(hi), wu));
                }
                const double tot = add_rn(imm, cont);
                if (d == 0 || tot > best) { best = tot; best_imm = imm; best_cons = cons; best_vol = vol; best_d = d; }
            };
            if (q.extra == 0) {   // the common case: withdraw, (nothing,) inject
                consider(0, w);
                consider(1, nd == 3 ? 0.0 : jn);
                if (nd == 3) consider(2, jn);
            } else {
                for (int d = 0; d < nd; d++) consider(d, volume_of(d));
            }
            const double next_inv = sub_rn(add_rn(v, best_vol), loss);
            q.inventory[n] = next_inv;
            q.pv_by_sim[n] = add_rn(q.pv_by_sim[n], best_imm);
            const double net = sub_rn(-best_vol, best_cons);
            sums[0] += v; sums[1] += best_vol; sums[2] += best_cons; sums[3] += loss; sums[4] += net; sums[5] += best_imm;
            sums[6] += mul_rn(-(add_rn(best_vol, best_cons)), s);
            if (q.o_inventory) __stcs(q.o_inventory + n, v);
            if (q.o_volume) __stcs(q.o_volume + n, best_vol);
            if (q.o_consumed) __stcs(q.o_consumed + n, best_cons);
            if (q.o_loss) __stcs(q.o_loss + n, loss);
            if (q.o_net) __stcs(q.o_net + n, net);
            if (q.o_pv) __stcs(q.o_pv + n, best_imm);
            if (q.o_decision) q.o_decision[n] = (int8_t)best_d;
        }
        // block reduction (fixed order)
#pragma unroll
        for (int i = 0; i < kFwdSums + KT; i++) {
            if (i < nsum) {
                double val = i < kFwdSums ? sums[i < kFwdSums ? i : 0] : xsum[i >= kFwdSums ? i - kFwdSums : 0];
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) val += __shfl_xor_sync(0xffffffffu, val, off);
                if (lane == 0) red[warp * nsum + i] = val;
            }
        }
        __syncthreads();
        for (int i = tid; i < nsum; i += blockDim.x) {
            double tot = 0.0;
            for (int w = 0; w < nwarps; w++) tot += red[w * nsum + i];
            q.sums[(size_t)blockIdx.x * nsum + i] = tot;
        }
    }
}

// -------------------------------------------------------------------------------------------------
// K4, straight-line form for the common shape (no extra decisions, K <= 16, fixed-spacing grids: checked on the host,
// lsmc_api.cu forward_fast_ok).  Same arithmetic as forward_kernel; the three bang-bang candidates are evaluated without a
// data-dependent branch (a one-sided set repeats its last candidate, which a strict '>' never selects), the grid lookup is
// locate_sm with selects instead of loops, both bracketing fitted values are always formed — so the compiler can
// interleave the candidates' dependent chains (grid loads -> division -> K-long FMA chains).
// -------------------------------------------------------------------------------------------------
#ifndef LSMC_FWDF_MINB
#define LSMC_FWDF_MINB 2
#endif
#ifndef LSMC_FWDF_THREADS
#define LSMC_FWDF_THREADS 256
#endif
constexpr int kFwdFastThreads = LSMC_FWDF_THREADS;

template <int KT>
__global__ void __launch_bounds__(kFwdFastThreads, LSMC_FWDF_MINB) forward_fast_kernel(const ForwardParams* __restrict__ pp, int p0, int p1, int max_grid) {
    extern __shared__ __align__(16) unsigned char fsm_fast[];   // same carve-up as forward_kernel (16-byte aligned beta rows)
    double* const fsm = reinterpret_cast<double*>(fsm_fast);
    __shared__ ForwardParams sq;
    __shared__ uint4 s_ops[kMaxOps];                 // basis program as element offsets into the thread's node column: dst, a, b
    __shared__ uint32_t s_basis_off[64];             // node column offset of basis function k
    __shared__ uint16_t s_class[(kFwdMaxGrid + 1) * 16];
    __shared__ double s_cx[kFwdPillars], s_clo[kFwdPillars], s_chi[kFwdPillars], s_slo[kFwdPillars], s_shi[kFwdPillars];
    constexpr int K = KT, BP = fwd_beta_pitch(KT), nsum = kFwdSums + K;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int nwarps = kFwdFastThreads / 32;
    constexpr int bdim = kFwdFastThreads;
    {
        const BasisDev* __restrict__ bd0 = pp[p0].basis;
        for (int o = tid; o < bd0->n_ops; o += bdim) {
            const BasisOpDev op = bd0->ops[o];
            s_ops[o] = make_uint4((uint32_t)op.dst * bdim, (uint32_t)op.a * bdim, (uint32_t)op.b * bdim, 0u);
        }
        for (int k = tid; k < 64; k += bdim) s_basis_off[k] = (uint32_t)bd0->node_of_basis[k] * bdim;
    }
    int class_n = -1;
    for (int p = p0; p < p1; p++) {
        __syncthreads();
        if (tid == 0) sq = pp[p];
        __syncthreads();
        const ForwardParams& q = sq;
        const BasisDev* __restrict__ bd = q.basis;
        const int n_ops = bd->n_ops, F = q.F;
        double* nodes = fsm + tid;
        double* beta_s = fsm + (size_t)bd->n_nodes * bdim;
        double* grid_s = beta_s + (size_t)max_grid * BP;
        double* red = grid_s + max_grid;
        const int Gn = q.Gn, top = Gn - 1;
        const bool fitted_mode = q.cont_mode == 0;
        {
            for (int l = warp; l < Gn; l += nwarps)
                for (int k = lane; k < K; k += 32)   // current period: the stored means as a constant-only "fit"
                    beta_s[l * BP + k] = fitted_mode ? __ldg(q.beta + (size_t)l * q.ldb + k) : (k == 0 ? __ldg(q.beta + (size_t)l * q.ldb) : 0.0);
            for (int l = tid; l < Gn; l += bdim) grid_s[l] = __ldg(q.next_grid + l);
            if (Gn != class_n)   // the table is a function of the number of grid points alone
                for (int e = tid; e < (Gn + 1) * 16; e += bdim) s_class[e] = fwd_class_result(Gn, (e >> 4) - 1, e & 15);
            class_n = Gn;
            if (q.c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR && q.c_n <= kFwdPillars && tid < kFwdPillars && tid >= q.c_n - 1)
                s_cx[tid] = __longlong_as_double(0x7ff0000000000000LL);   // interior pillars end at c_n - 2: the count below stops there
            if (q.c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR && q.c_n <= kFwdPillars && tid < q.c_n) {
                const double x0 = __ldg(q.c_x + tid), l0 = __ldg(q.c_lo + tid), h0 = __ldg(q.c_hi + tid);
                if (tid < q.c_n - 1) s_cx[tid] = x0;
                s_clo[tid] = l0; s_chi[tid] = h0;
                if (tid + 1 < q.c_n) {   // slopes as spline_dev forms them
                    const double dx = __ldg(q.c_x + tid + 1) - x0;
                    s_slo[tid] = (__ldg(q.c_lo + tid + 1) - l0) / dx;
                    s_shi[tid] = (__ldg(q.c_hi + tid + 1) - h0) / dx;
                }
            }
        }
        __syncthreads();
        const bool staged_pillars = q.c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR && q.c_n <= kFwdPillars && q.c_n >= 2;
        const double g0 = grid_s[0], inv_h = Gn > 1 ? 1.0 / (grid_s[1] - grid_s[0]) : 0.0;
        const double g_top = grid_s[top], g_pre = grid_s[max(top - 1, 0)];
        // period constants in registers
        const double df = q.df, tol = q.tol, loss_pct = q.loss_pct, next_min = q.next_min, next_max = q.next_max;
        const double cons_inj = q.cons_inj, cons_wd = q.cons_wd, inj_cost = q.inj_cost, wd_cost = q.wd_cost, inj_df = q.inj_df, wd_df = q.wd_df;
        const double inv_cost = q.inv_cost, inv_cost_df = q.inv_cost_df;
        const bool spot_is_const = q.spot_is_const != 0;
        const double spot_const = q.spot_const;
        const double* __restrict__ spot_row = q.spot;
        const double* __restrict__ fac_rows = q.fac;
        double* __restrict__ inv_state = q.inventory;
        double* __restrict__ pv_state = q.pv_by_sim;
        const int64_t N = q.N, Npad = q.Npad;
        const int c_kind = q.c_kind, c_n = q.c_n;
        const bool any_panel = q.o_inventory || q.o_volume || q.o_consumed || q.o_loss || q.o_net || q.o_pv || q.o_decision;

        double sums[kFwdSums], xsum[KT];
#pragma unroll
        for (int i = 0; i < kFwdSums; i++) sums[i] = 0.0;
#pragma unroll
        for (int k = 0; k < KT; k++) xsum[k] = 0.0;

        const int64_t stride = (int64_t)gridDim.x * bdim;
        for (int64_t n = (int64_t)blockIdx.x * bdim + tid; n < N; n += stride) {
            const double s = spot_is_const ? spot_const : __ldcs(spot_row + n);
            const double v = inv_state[n];
            const double pv_old = pv_state[n];
            double xrow[KT];
            if (fitted_mode) {
                nodes[0] = 0.0; nodes[bdim] = 1.0; nodes[2 * bdim] = s;
                {
                    const double* fp = fac_rows + n;
                    double* np = nodes + 3 * bdim;
                    for (int f = 0; f < F; f++, fp += Npad, np += bdim) *np = __ldcs(fp);
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
            if (staged_pillars) {   // spline_dev on the staged table: segment by counting, slopes precomputed
                int k = 0;
#pragma unroll
                for (int i = 1; i <= 6; i++) k += (s_cx[i] <= v) ? 1 : 0;   // entries from c_n - 1 on are +inf
                for (int i = 7; i < c_n - 1; i++) k += (s_cx[i] <= v) ? 1 : 0;
                const double dxv = sub_rn(v, s_cx[k]);
                rmax = add_rn(s_chi[k], mul_rn(dxv, s_shi[k]));
                rmin = add_rn(s_clo[k], mul_rn(dxv, s_slo[k]));
            } else if (c_kind == LSMC_CONSTRAINT_CONSTANT) { rmin = __ldg(q.c_lo); rmax = __ldg(q.c_hi); }
            else if (c_kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) { rmax = spline_dev(q.c_x, q.c_hi, c_n, v); rmin = spline_dev(q.c_x, q.c_lo, c_n, v); }
            else if (c_kind == LSMC_CONSTRAINT_POLYNOMIAL) { rmax = horner_dev(q.c_hi, c_n, v); rmin = horner_dev(q.c_lo, c_n, v); }
            else {
                if (v < __ldg(q.c_x) - q.slack_hi || v > __ldg(q.c_x + c_n - 1) + q.slack_hi) atomicExch(q.error_flag, 2);
                int idx = min(max(last_at_or_below(q.c_x, c_n, v), 0), c_n - 1);
                rmin = __ldg(q.c_lo + idx); rmax = __ldg(q.c_hi + idx);
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
            const double inv_cost_npv = inv_cost == 0.0 ? 0.0 : mul_rn(mul_rn(v, inv_cost), inv_cost_df);
            double vol[3], tot[3], imm[3], cons[3];
            vol[0] = w; vol[1] = one_sided ? jn : 0.0; vol[2] = jn;
            bool all_found = true;
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double vd = vol[d];
                const double after = sub_rn(add_rn(v, vd), loss);
                const bool pos = vd > 0.0;
                const double avol = fabs(vd);
                cons[d] = mul_rn(pos ? cons_inj : cons_wd, avol);
                const double a = mul_rn(mul_rn(-vd, s), df);
                const double c = mul_rn(mul_rn(-cons[d], s), df);
                const double cost = mul_rn(mul_rn(pos ? inj_cost : wd_cost, avol), pos ? inj_df : wd_df);
                imm[d] = sub_rn(add_rn(sub_rn(a, cost), c), inv_cost_npv);
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
                const double* bl = beta_s + lo * BP;
                const double* bh = beta_s + hi * BP;
                double f_lo = 0.0, f_hi = 0.0;
#pragma unroll
                for (int k = 0; k < KT; k++) { f_lo = fma(xrow[k], bl[k], f_lo); f_hi = fma(xrow[k], bh[k], f_hi); }
                const double wl = sub_rn(gu, after) / sub_rn(gu, gl);
                const double wu = sub_rn(1.0, wl);
                const double cont = lo == hi ? f_lo : add_rn(mul_rn(f_lo, wl), mul_rn(f_hi, wu));
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
        // block reduction (fixed order)
#pragma unroll
        for (int i = 0; i < kFwdSums + KT; i++) {
            double val = i < kFwdSums ? sums[i < kFwdSums ? i : 0] : xsum[i >= kFwdSums ? i - kFwdSums : 0];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) val += __shfl_xor_sync(0xffffffffu, val, off);
            if (lane == 0) red[warp * nsum + i] = val;
        }
        __syncthreads();
        for (int i = tid; i < nsum; i += bdim) {
            double totv = 0.0;
            for (int w = 0; w < nwarps; w++) totv += red[w * nsum + i];
            q.sums[(size_t)blockIdx.x * nsum + i] = totv;
        }
    }
}

template <int KT>
static void launch_forward_fast_t(const ForwardParams* pp, int p0, int p1, int n_nodes, int max_grid, int blocks, cudaStream_t stream) {
    const size_t smem = sizeof(double) * ((size_t)n_nodes * kFwdFastThreads + (size_t)max_grid * (fwd_beta_pitch(KT) + 1) + (size_t)(kFwdFastThreads / 32) * (kFwdSums + KT));
    if (ensure_dynamic_smem((const void*)forward_fast_kernel<KT>, smem) != cudaSuccess) return;
    forward_fast_kernel<KT><<<blocks, kFwdFastThreads, smem, stream>>>(pp, p0, p1, max_grid);
}

// Dynamic shared memory the forward pass needs for this shape (checked by lsmc_set_problem before anything runs).
size_t forward_smem_bytes(int K, int n_nodes, int max_grid, bool fast) {
    const int KT = (fast || K <= 16) ? K : 64, threads = fast ? kFwdFastThreads : kFwdThreads;
    return sizeof(double) * ((size_t)n_nodes * threads + (size_t)max_grid * (fwd_beta_pitch(KT) + 1) + (size_t)(threads / 32) * (kFwdSums + K));
}

int forward_fast_blocks_per_sm() { return LSMC_FWDF_MINB; }
int forward_fast_threads() { return kFwdFastThreads; }

template <int KT, bool EXACT>
static void launch_forward_t(const ForwardParams* pp, int p0, int p1, int K, int n_nodes, int max_grid, int blocks, cudaStream_t stream) {
    const size_t smem = sizeof(double) * ((size_t)n_nodes * kFwdThreads + (size_t)max_grid * (fwd_beta_pitch(KT) + 1) + (size_t)(kFwdThreads / 32) * (kFwdSums + K));
    if (ensure_dynamic_smem((const void*)forward_kernel<KT, EXACT>, smem) != cudaSuccess) return;
    forward_kernel<KT, EXACT><<<blocks, kFwdThreads, smem, stream>>>(pp, p0, p1, max_grid);
}

void launch_forward(const ForwardParams* pp, int p0, int p1, int K, int n_nodes, int max_grid, int blocks, bool fast, cudaStream_t stream) {
    if (fast) {
        switch (K) {
#define FAST_K(k) case k: launch_forward_fast_t<k>(pp, p0, p1, n_nodes, max_grid, blocks, stream); return;
#ifdef LSMC_FAST_BUILD
            FAST_K(3) FAST_K(10)
#else
            FAST_K(1) FAST_K(2) FAST_K(3) FAST_K(4) FAST_K(5) FAST_K(6) FAST_K(7) FAST_K(8) FAST_K(9) FAST_K(10)
            FAST_K(11) FAST_K(12) FAST_K(13) FAST_K(14) FAST_K(15) FAST_K(16)
#endif
#undef FAST_K
            default: break;
        }
    }
    switch (K) {
#define EXACT_K(k) case k: launch_forward_t<k, true>(pp, p0, p1, K, n_nodes, max_grid, blocks, stream); return;
#ifdef LSMC_FAST_BUILD
        EXACT_K(3) EXACT_K(10)
#else
        EXACT_K(1) EXACT_K(2) EXACT_K(3) EXACT_K(4) EXACT_K(5) EXACT_K(6) EXACT_K(7) EXACT_K(8) EXACT_K(9) EXACT_K(10)
        EXACT_K(11) EXACT_K(12) EXACT_K(13) EXACT_K(14) EXACT_K(15) EXACT_K(16)
#endif
#undef EXACT_K
#ifndef LSMC_FAST_BUILD
        default: launch_forward_t<64, false>(pp, p0, p1, K, n_nodes, max_grid, blocks, stream); return;
#else
        default: return;
#endif
    }
}

// Column reduction of block partials: out[r][c] = sum_b in[r][b][c]; one warp per (r, c).
__global__ void __launch_bounds__(256) reduce_rows_kernel(const double* __restrict__ in, int n_rows, int n_blocks, int n_cols,
                                                          double* __restrict__ out) {
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (e >= n_rows * n_cols) return;
    const int r = e / n_cols, c = e % n_cols;
    double s = 0.0;
    for (int b = lane; b < n_blocks; b += 32) s += in[((size_t)r * n_blocks + b) * n_cols + c];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) out[e] = s;
}

void launch_reduce_rows(const double* in, int n_rows, int n_blocks, int n_cols, double* out, cudaStream_t stream) {
    const int wpb = 8, n = n_rows * n_cols;
    reduce_rows_kernel<<<(n + wpb - 1) / wpb, wpb * 32, 0, stream>>>(in, n_rows, n_blocks, n_cols, out);
}

// Final period (LSV.cs:563-579) + sums for NPV / final inventory / standard error.
//   pass 0: pv += terminal value on the regression sims' end prices; block sums {final inventory, terminal pv, pv}
//   pass 1: sum of squared deviations of (pair means of) pv from the mean
__global__ void __launch_bounds__(256) final_kernel(FinalParams q) {
    __shared__ double sm[8][3];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (q.pass == 0) {
        for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < q.N; n += stride) {
            const double inv = q.inventory[n];
            double tpv = 0.0;
            if (!q.must_be_empty) {
                const double s = q.spot_end[n];
                tpv = add_rn(add_rn(add_rn(q.tc0, mul_rn(q.tc1, s)), mul_rn(q.tc2, inv)), mul_rn(q.tc3, mul_rn(s, inv)));
                q.pv_by_sim[n] = add_rn(q.pv_by_sim[n], tpv);
            }
            if (q.o_inventory) q.o_inventory[n] = inv;
            if (q.o_pv) q.o_pv[n] = tpv;
            a0 += inv; a1 += tpv; a2 += q.pv_by_sim[n];
        }
    } else {
        const int64_t m = q.paired ? q.N / 2 : q.N;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
            const double v = q.paired ? 0.5 * (q.pv_by_sim[2 * i] + q.pv_by_sim[2 * i + 1]) : q.pv_by_sim[i];
            const double dlt = v - q.mean;
            a0 += dlt * dlt;
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        a0 += __shfl_xor_sync(0xffffffffu, a0, off);
        a1 += __shfl_xor_sync(0xffffffffu, a1, off);
        a2 += __shfl_xor_sync(0xffffffffu, a2, off);
    }
    if (lane == 0) { sm[warp][0] = a0; sm[warp][1] = a1; sm[warp][2] = a2; }
    __syncthreads();
    if (threadIdx.x < 3) {
        double tot = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) tot += sm[w][threadIdx.x];
        q.sums[(size_t)blockIdx.x * 3 + threadIdx.x] = tot;
    }
}

void launch_final(const FinalParams& q, int blocks, cudaStream_t stream) { final_kernel<<<blocks, 256, 0, stream>>>(q); }

__global__ void fill_kernel(double* p, int64_t n, double v) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

void launch_fill(double* p, int64_t n, double v, cudaStream_t stream) {
    int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
    fill_kernel<<<blocks, 256, 0, stream>>>(p, n, v);
}

}  // namespace lsmc
