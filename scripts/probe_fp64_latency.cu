// probe_fp64_latency.cu — FP64 pipe throughput of a B200 SM as a function of resident warps and independent chains per
// warp (DMUL+DADD pairs, the unfused arithmetic of the decision loop), and DMMA 8x8x4 issue rate from few warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_lat scripts/probe_fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int CH, int MODE>
__global__ void probe(double* out, int iters, double seed) {
    double a = seed + threadIdx.x * 1e-9, b = 1.0 - 1e-12;
    double acc[CH], m[2 * CH];
#pragma unroll
    for (int i = 0; i < CH; i++) { acc[i] = i * 1e-3; m[2 * i] = m[2 * i + 1] = 0.0; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < CH; i++) {
                if (MODE == 0) acc[i] = __dadd_rn(__dmul_rn(acc[i], b), a);
                else dmma884(m[2 * i], m[2 * i + 1], a, b);
            }
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < CH; i++) s += acc[i] + m[2 * i] + m[2 * i + 1];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CH, int MODE>
void run(int threads, int sms) {
    const int iters = 5000;
    double* out;
    cudaMalloc(&out, sizeof(double) * threads * sms);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<CH, MODE><<<sms, threads>>>(out, 100, 0.5);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    probe<CH, MODE><<<sms, threads>>>(out, iters, 0.5);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double inst_per_warp = (double)iters * 4 * CH * (MODE == 0 ? 2 : 1);
    const double clk = ms * 1e-3 * 1.965e9;
    const double warps_per_smsp = threads / 32 / 4.0;
    printf("%s warps/SMSP %4.1f chains %d : %7.3f ms  clk per %s per SMSP %6.2f  (per-warp dependent step %6.1f clk)\n", MODE == 0 ? "DMUL+DADD" : "DMMA     ",
           warps_per_smsp, CH, ms, MODE == 0 ? "FP64 inst" : "DMMA", clk / (inst_per_warp * warps_per_smsp), clk / (iters * 4.0 * (MODE == 0 ? 2 : 1)));
    cudaFree(out);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    for (int threads : {128, 256, 512, 1024}) {
        run<1, 0>(threads, sms); run<2, 0>(threads, sms); run<4, 0>(threads, sms); run<6, 0>(threads, sms); run<8, 0>(threads, sms);
    }
    for (int threads : {128, 256, 512}) { run<1, 1>(threads, sms); run<2, 1>(threads, sms); run<4, 1>(threads, sms); run<6, 1>(threads, sms); }
    return 0;
}
