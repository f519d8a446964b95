// probe_fp64.cu — measures the FP64 roofs of a B200 that the sweep kernel is bounded by besides HBM:
// scalar DFMA/DMUL/DADD issue rate, DMMA (mma.sync m8n8k4 / m16n8k8 f64) rate, and both interleaved.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_fp64 scripts/probe_fp64.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double& d0, double& d1, double& d2, double& d3, double a0, double a1, double a2, double a3, double b0, double b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+d"(d0), "+d"(d1), "+d"(d2), "+d"(d3) : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1));
}

template <int MODE>
__global__ void __launch_bounds__(512) probe(double* out, int iters, double seed) {
    double a = seed + threadIdx.x * 1e-9, b = 1.0 - 1e-12;
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = i * 1e-3;
    double m[8];
#pragma unroll
    for (int i = 0; i < 8; i++) m[i] = 0.0;
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {   // 8 independent DFMA chains
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = fma(acc[i], b, a);
        } else if (MODE == 1) {   // DMUL + DADD (unfused, like the reference arithmetic)
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = __dadd_rn(__dmul_rn(acc[i], b), a);
        } else if (MODE == 2) {   // 4 independent DMMA 8x8x4 chains
#pragma unroll
            for (int i = 0; i < 4; i++) dmma884(m[2 * i], m[2 * i + 1], a, b);
        } else if (MODE == 3) {   // 2 independent DMMA 16x8x8 chains
#pragma unroll
            for (int i = 0; i < 2; i++) dmma1688(m[4 * i], m[4 * i + 1], m[4 * i + 2], m[4 * i + 3], a, b, a, b, b, a);
        } else if (MODE == 4) {   // interleaved: 4 DMMA + 8 DFMA per iteration
#pragma unroll
            for (int i = 0; i < 4; i++) {
                dmma884(m[2 * i], m[2 * i + 1], a, b);
                acc[2 * i] = fma(acc[2 * i], b, a);
                acc[2 * i + 1] = fma(acc[2 * i + 1], b, a);
            }
        } else if (MODE == 5) {   // DSETP + select (compare path)
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = (acc[i] > a) ? b : acc[i] + 1e-9;
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i] + m[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, double ops_per_thread_iter, double flops_per_thread_iter, int sms) {
    const int threads = 512, blocks = sms * 2, iters = 20000;
    double* out;
    cudaMalloc(&out, sizeof(double) * threads * blocks);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<blocks, threads>>>(out, 1000, 0.5);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    probe<MODE><<<blocks, threads>>>(out, iters, 0.5);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double total_threads = (double)threads * blocks;
    const double warp_inst = total_threads / 32 * iters * ops_per_thread_iter;
    const double tflops = total_threads * iters * flops_per_thread_iter / (ms * 1e-3) / 1e12;
    printf("%-34s %8.3f ms  %7.2f TFLOP/s  warp-inst/SM/us %8.1f  (err %s)\n", name, ms, tflops, warp_inst / sms / (ms * 1e3),
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%s, %d SMs, max clock %d MHz\n", p.name, p.multiProcessorCount, clk / 1000);
    const int sms = p.multiProcessorCount;
    run<0>("DFMA x8 chains", 8, 16, sms);
    run<1>("DMUL+DADD x8 chains", 16, 16, sms);
    run<2>("DMMA m8n8k4 x4 chains", 4, 4 * 2.0 * 8 * 8 * 4 / 32, sms);
    run<3>("DMMA m16n8k8 x2 chains", 2, 2 * 2.0 * 16 * 8 * 8 / 32, sms);
    run<4>("4 DMMA m8n8k4 + 8 DFMA interleaved", 12, 4 * 2.0 * 8 * 8 * 4 / 32 + 16, sms);
    run<5>("DSETP+select+DADD x8", 24, 8, sms);
    return 0;
}
