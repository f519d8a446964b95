// Development probe: shared-memory pipe cost (clk per warp-level load at full occupancy) of the access patterns of the
// sweep's decision loop.  nvcc -arch=sm_100a -O3 -o probe_lds probe_lds.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int VEC, int PAT>
__global__ void k(double* out, int iters, long long* clk) {
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i;
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    int off;   // element offset
    if (PAT == 0) off = 0;                               // all lanes one address
    else if (PAT == 1) off = VEC * t;                    // 4 distinct addresses (by t), contiguous
    else if (PAT == 2) off = VEC * lane;                 // fully distinct contiguous
    else if (PAT == 3) off = 56 * g + VEC * t;           // 8 rows of pitch 448 B, 4 x VEC per row (C^ pattern)
    else if (PAT == 4) off = 60 * g + 2 * t;             // V gather pattern: pitch 480 B
    else off = VEC * (lane & 7);                         // 8 distinct
    unsigned acc = 0;
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm + off);
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const unsigned a = base + 8u * ((u * 64) & 1023);
            if (VEC == 2) { unsigned long long x, y; asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "r"(a)); acc ^= (unsigned)x ^ (unsigned)(x >> 32) ^ (unsigned)y ^ (unsigned)(y >> 32); }
            else { unsigned long long x; asm volatile("ld.shared.u64 %0, [%1];" : "=l"(x) : "r"(a)); acc ^= (unsigned)x ^ (unsigned)(x >> 32); }
        }
    }
    long long t1 = clock64();
    if (acc == 12345u) out[0] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
template <int VEC, int PAT>
void run(const char* name) {
    double* out; long long* clk; cudaMalloc(&out, 8); cudaMalloc(&clk, 8);
    const int iters = 2000;
    k<VEC, PAT><<<148, 512, 40000>>>(out, iters, clk);
    k<VEC, PAT><<<148, 512, 40000>>>(out, iters, clk);
    long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    // 16 warps per SM each issuing iters*16 loads: SM-level clk per warp load
    printf("%-44s clk per warp-load (SM level) %.2f   err %s\n", name, (double)h / (iters * 16.0 * 16.0), cudaGetErrorString(cudaGetLastError()));
}
int main() {
    run<2, 0>("LDS.128 all lanes same address");
    run<2, 1>("LDS.128 4 addresses by t (64 B)");
    run<2, 5>("LDS.128 8 addresses (128 B)");
    run<2, 2>("LDS.128 32 distinct (512 B)");
    run<2, 3>("LDS.128 C^ pattern (8 rows pitch 448)");
    run<1, 0>("LDS.64 all lanes same address");
    run<1, 1>("LDS.64 4 addresses by t (32 B)");
    run<1, 2>("LDS.64 32 distinct (256 B)");
    run<1, 4>("LDS.64 gather pattern (8 rows pitch 480)");
    return 0;
}
