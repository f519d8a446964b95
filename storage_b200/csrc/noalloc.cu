// noalloc.cu — lsmc_npv_stats_noalloc: the NPV / standard-error reduction of the forward pass (kernels.cu final_kernel +
// reduce_rows_kernel: LSV.cs:584 and the antithetic pair-mean standard error) for callers that cannot allocate device
// memory: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// It has no counterpart in the reference.
//
// The block partials that final_kernel writes to a cudaMalloc'ed buffer live here in a file-scope __device__ array of a
// fixed, modest size (every GPU that loads the library holds it until exit).  The entry point calls no cudaMalloc, cudaFree
// or allocating library routine.  In host code the array's NAME is a host-side stand-in, not a device address: the
// address the kernels receive is obtained once per device with cudaGetSymbolAddress while that device is current.
//
// Every call shares the array and finds the previous call's contents in it.  Nothing has to start at zero: a call writes
// every slot it later reads (chunk sums -> chunk squared deviations -> results) before it reads it.  Two calls that
// overlap on one GPU (different streams, or different host threads) race on the array: serialise them.
#include <cstdio>
#include <cuda_runtime.h>

#include "../../include/lsmc_b200.h"

namespace {

constexpr int kChunk = 8192;                                   // paths per CTA: a fixed partition, so the sums do not depend on the GPU
constexpr int kMaxChunks = LSMC_NOALLOC_MAX_SIMS / kChunk;     // 2048
constexpr int kThreads = 256;
constexpr int kMaxDevices = 64;

// [0, kMaxChunks): chunk sums of pv; [kMaxChunks, 2 kMaxChunks): chunk sums of squared deviations.  32 KB.
static __device__ __align__(256) double g_noalloc_scratch[2 * kMaxChunks];

__device__ __forceinline__ double block_sum_256(double v, double* sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    __syncthreads();   // sh may still be read by the previous reduction
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; w++) tot += sh[w];
    return tot;
}

// Sum of all chunk sums in a fixed order (thread t takes chunks t, t + 256, ...): every CTA that calls it gets the same bits.
__device__ __forceinline__ double total_of_chunks(const double* __restrict__ sums, int n_chunks, double* sh) {
    double v = 0.0;
    for (int c = threadIdx.x; c < n_chunks; c += kThreads) v += sums[c];
    return block_sum_256(v, sh);
}

// pass 0: scratch[chunk] = sum of pv over the chunk.
__global__ void __launch_bounds__(kThreads) noalloc_sum_kernel(const double* __restrict__ pv, long long n, double* __restrict__ scratch) {
    __shared__ double sh[kThreads / 32];
    const long long base = (long long)blockIdx.x * kChunk;
    double v = 0.0;
    for (int i = threadIdx.x; i < kChunk; i += kThreads) {
        const long long idx = base + i;
        if (idx < n) v += pv[idx];
    }
    const double tot = block_sum_256(v, sh);
    if (threadIdx.x == 0) scratch[blockIdx.x] = tot;
}

// pass 1: scratch[kMaxChunks + chunk] = sum over the chunk of (x - mean)^2, x = pv or the mean of an antithetic pair.
__global__ void __launch_bounds__(kThreads) noalloc_dev_kernel(const double* __restrict__ pv, long long n, int paired, int n_chunks,
                                                               double* __restrict__ scratch) {
    __shared__ double sh[kThreads / 32];
    const double mean = total_of_chunks(scratch, n_chunks, sh) / (double)n;
    const long long m = paired ? n / 2 : n;
    const int per = paired ? kChunk / 2 : kChunk;               // a chunk of 8192 paths holds 4096 pairs
    const long long base = (long long)blockIdx.x * per;
    double v = 0.0;
    for (int i = threadIdx.x; i < per; i += kThreads) {
        const long long idx = base + i;
        if (idx < m) {
            const double x = paired ? 0.5 * (pv[2 * idx] + pv[2 * idx + 1]) : pv[idx];
            const double d = x - mean;
            v += d * d;
        }
    }
    const double tot = block_sum_256(v, sh);
    if (threadIdx.x == 0) scratch[kMaxChunks + blockIdx.x] = tot;
}

// pass 2: out = {npv, standard error, sum of pv, sum of squared deviations}.
__global__ void __launch_bounds__(kThreads) noalloc_final_kernel(long long n, int paired, int n_chunks, const double* __restrict__ scratch,
                                                                 double* __restrict__ out) {
    __shared__ double sh[kThreads / 32];
    const double sum = total_of_chunks(scratch, n_chunks, sh);
    const double ss = total_of_chunks(scratch + kMaxChunks, n_chunks, sh);
    if (threadIdx.x == 0) {
        const double m = (double)(paired ? n / 2 : n);
        out[0] = sum / (double)n;
        out[1] = m > 1.0 ? sqrt(ss / (m - 1.0) / m) : 0.0;
        out[2] = sum;
        out[3] = ss;
    }
}

int fail(char* err, size_t err_len, int status, const char* what, const char* detail) {
    if (err && err_len) snprintf(err, err_len, "%s%s%s", what, detail ? ": " : "", detail ? detail : "");
    return status;
}

}  // namespace

extern "C" int lsmc_npv_stats_noalloc(const double* d_pv_by_sim, int64_t n_sims, int antithetic_pairs, double* d_out4,
                                      void* cuda_stream, char* err, size_t err_len) {
    if (err && err_len) err[0] = '\0';
    if (!d_pv_by_sim || !d_out4) return fail(err, err_len, LSMC_EINVAL, "lsmc_npv_stats_noalloc: d_pv_by_sim and d_out4 must be device pointers, not NULL", nullptr);
    if (n_sims < 1) return fail(err, err_len, LSMC_EINVAL, "lsmc_npv_stats_noalloc: n_sims must be positive", nullptr);
    if (n_sims > (int64_t)LSMC_NOALLOC_MAX_SIMS) {
        char msg[160];
        snprintf(msg, sizeof msg, "n_sims = %lld exceeds LSMC_NOALLOC_MAX_SIMS = %d (the static scratch array holds one partial per %d paths)",
                 (long long)n_sims, (int)LSMC_NOALLOC_MAX_SIMS, kChunk);
        return fail(err, err_len, LSMC_EUNSUPPORTED, "lsmc_npv_stats_noalloc", msg);
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(err, err_len, LSMC_ECUDA, "cudaGetDevice", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail(err, err_len, LSMC_EUNSUPPORTED, "lsmc_npv_stats_noalloc: device index above 63", nullptr);
    // set-up, once per device, with that device current: the device address of the scratch array
    static double* scratch_of_device[kMaxDevices];   // zero-initialised; a benign race stores the same value twice
    double* scratch = scratch_of_device[dev];
    if (!scratch) {
        e = cudaGetSymbolAddress((void**)&scratch, g_noalloc_scratch);
        if (e != cudaSuccess || !scratch) return fail(err, err_len, LSMC_ECUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", cudaGetErrorString(e));
        scratch_of_device[dev] = scratch;
    }
    const int paired = (antithetic_pairs && n_sims % 2 == 0) ? 1 : 0;   // lsmc_api.cu forward(): pair means only when every path has its partner
    const int n_chunks = (int)((n_sims + kChunk - 1) / kChunk);
    cudaStream_t s = (cudaStream_t)cuda_stream;
    noalloc_sum_kernel<<<n_chunks, kThreads, 0, s>>>(d_pv_by_sim, (long long)n_sims, scratch);
    noalloc_dev_kernel<<<n_chunks, kThreads, 0, s>>>(d_pv_by_sim, (long long)n_sims, paired, n_chunks, scratch);
    noalloc_final_kernel<<<1, kThreads, 0, s>>>((long long)n_sims, paired, n_chunks, scratch, d_out4);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(err, err_len, LSMC_ECUDA, "lsmc_npv_stats_noalloc launch", cudaGetErrorString(e));
    return LSMC_OK;
}
