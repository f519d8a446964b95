// tail.cu — what follows the sweep of one period, as ONE kernel: the deterministic reduction of the sweep's per-CTA
// normal-equation partials, the exchange of the reduced block between the ranks over peer memory (peer.h), and the
// solve of the K x K system for every grid point (solve_device.cuh), spread over the CTAs of the grid.
//
// It replaces reduce_partials_kernel -> ncclAllReduce -> solve_kernel (three launches, the last on a single CTA).
// Phases, for a grid of NC co-resident CTAs of 128 threads (NC <= number of SMs, so the spin below cannot starve):
//   1  element e of the block is reduced by one warp (lanes stride over the partials, shuffle tree: the same summation
//      order as reduce_partials_kernel) and stored into slot [parity][rank] of every rank's exchange buffer;
//   2  arrive at every rank's counter (one system-scope atomic per CTA and rank) and wait until this rank's own counter
//      shows that every CTA of every rank has arrived — with one rank this is simply the grid barrier between 1 and 3;
//   3  every CTA factorises the (rank-order summed) Gram matrix redundantly and solves the right-hand sides
//      blockIdx.x, blockIdx.x + NC, ...  (K <= 16: about a microsecond; no second barrier, no broadcast of the factor).
#include "peer.h"
#include "solve_device.cuh"

namespace lsmc {

namespace {

constexpr int kTailThreads = 128;

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// One thread per CTA: publish this CTA's slot stores, arrive everywhere, wait for everybody.  With one rank this is a plain
// grid barrier: device scope is enough (no system-scope fence, no system-scope atomic: about 2 us less per period).  With
// several ranks every CTA fences its peer stores at system scope and takes a device-scope ticket; the CTA that draws the last
// ticket of the exchange (fence - atomic - fence: it has observed all the others' fences) adds the rank's CTA count to every
// rank's counter.
__device__ __forceinline__ void arrive_and_wait(const PeerView& pv, unsigned long long target, unsigned long long local_target, int* error_flag) {
    __syncthreads();   // every thread's slot stores happen-before the fence below
    if (threadIdx.x == 0) {
        const unsigned long long* mine = pv.counter[pv.rank];
        const long long t0 = clock64();
        if (pv.world == 1) {
            __threadfence();
            atomicAdd(pv.counter[0], 1ULL);
            while (ld_acquire_gpu(mine) < target) {
                if (clock64() - t0 > 4000000000LL) { atomicExch(error_flag, kPeerTimeout); break; }
            }
        } else {
            __threadfence_system();
            if (atomicAdd(pv.ticket, 1ULL) + 1ULL == local_target) {
                __threadfence_system();
                for (int r = 0; r < pv.world; r++) atomicAdd_system(pv.counter[r], (unsigned long long)gridDim.x);
            }
            // a peer that failed or was cancelled never arrives: give up after ~2 s, and at once in every later exchange
            while (ld_acquire_sys(mine) < target) {
                if (*reinterpret_cast<volatile int*>(error_flag) == kPeerTimeout) break;
                if (clock64() - t0 > 4000000000LL) { atomicExch(error_flag, kPeerTimeout); break; }
                __nanosleep(20);
            }
        }
    }
    __syncthreads();
}

// The exchanged block: sum over the ranks' slots, in rank order, bypassing L1 (the lines were written by peers).
struct SlotRed {
    const double* __restrict__ base;   // this rank's slots of the current parity: [world][kPeerSlotDoubles]
    int world, ld;
    // Eight predicated loads in flight, then the sum in rank order (an absent rank adds +0.0: the value is unchanged).  A loop over
    // `world` was unrolled with a remainder at each of the thirty-odd places this is inlined: 2 600 of the kernel's 7 500
    // instructions — and the tail kernel runs its code once per launch, from an instruction cache the sweep has just emptied.
    __device__ __forceinline__ double at(int e) const {
        const double* p = base + e;
        double v[kMaxPeers];
#pragma unroll
        for (int r = 0; r < kMaxPeers; r++) v[r] = r < world ? __ldcg(p + (size_t)r * kPeerSlotDoubles) : 0.0;
        double s = v[0];
#pragma unroll
        for (int r = 1; r < kMaxPeers; r++) s += v[r];
        return s;
    }
    __device__ __forceinline__ double operator()(int k, int col) const { return at(k * ld + col); }
};

__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

template <int R>   // rows of the system per lane: 1 (K <= 32) or 2 (K <= 64)
__global__ void __launch_bounds__(kTailThreads) tail_kernel(const TailParams tp) {
    extern __shared__ double tsm[];
    __shared__ int dropped[64];
    pdl_launch_dependents();   // PDL: the next sweep's CTAs set themselves up while this kernel runs ...
    if (tp.do_solve) solve_stage_precond(tp.solve, tsm);   // (per-problem constant, written before the first sweep: safe ahead of the wait)
    pdl_wait();                // ... and this one starts the moment the sweep before it has completed
    long long* tr = (tp.trace && blockIdx.x == 0 && threadIdx.x == 0) ? tp.trace : nullptr;
    if (tr) tr[0] = gtimer();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = kTailThreads / 32;
    const PeerView& pv = tp.peers;
    const size_t par_off = (size_t)tp.parity * pv.world * kPeerSlotDoubles;
    // ---- 1: reduce the partials straight into every rank's slot of this rank.  Eight lanes per element: lane j of the
    //         group sums partials j, j + 8, ... (eight loads in flight per lane), then a three-step shuffle tree: a fixed
    //         order, the same on every run and every rank count ----
    {
        const int sub = lane & 7, grp = lane >> 3;
        for (int e0 = (blockIdx.x * wpb + warp) * 4; e0 < tp.n_elem; e0 += gridDim.x * wpb * 4) {
            const int e = e0 + grp;
            double s = 0.0;
            if (e < tp.n_elem) {
                const double* src = tp.partial + e;
                int c = sub;
                for (; c + 56 < tp.n_parts; c += 64) {
                    double v[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) v[u] = src[(size_t)(c + 8 * u) * tp.n_elem];
#pragma unroll
                    for (int u = 0; u < 8; u++) s += v[u];
                }
                for (; c < tp.n_parts; c += 8) s += src[(size_t)c * tp.n_elem];
            }
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            if (e < tp.n_elem && sub < pv.world) pv.slots[sub][par_off + (size_t)pv.rank * kPeerSlotDoubles + e] = s;
        }
    }
    // ---- 2: exchange ----
    if (tr) tr[1] = gtimer();
    arrive_and_wait(pv, tp.target, tp.local_target, tp.error_flag);
    if (tr) tr[2] = gtimer();
    const SlotRed red{pv.slots[pv.rank] + par_off, pv.world, tp.solve.ncols};
    if (blockIdx.x == 0 && (int)threadIdx.x < tp.n_out) tp.total_out[threadIdx.x] = red.at(threadIdx.x);
    // ---- 3: solve ----
    // (A register-resident variant — rows of L, L^T, T per lane, fully unrolled for K <= 16 — was measured and dropped: 34 us
    //  against 21 us; a kernel that runs its straight-line code once per launch pays for every instruction fetch.)
    if (tp.do_solve) solve_cta_warps<R>(tp.solve, red, tsm, dropped, blockIdx.x, gridDim.x, tr, true);
    if (tr) tr[6] = gtimer();
}

__global__ void __launch_bounds__(256) peer_small_kernel(const PeerSmallParams pp) {
    const PeerView& pv = pp.peers;
    const size_t par_off = (size_t)pp.parity * pv.world * kPeerSlotDoubles;
    const bool bcast = pp.root >= 0;
    if (!bcast || pv.rank == pp.root)
        for (int i = threadIdx.x; i < pp.count; i += blockDim.x) {
            const double v = pp.buf[i];
            for (int r = 0; r < pv.world; r++) pv.slots[r][par_off + (size_t)pv.rank * kPeerSlotDoubles + i] = v;
        }
    arrive_and_wait(pv, pp.target, pp.local_target, pp.error_flag);
    const double* mine = pv.slots[pv.rank] + par_off;
    for (int i = threadIdx.x; i < pp.count; i += blockDim.x) {
        double s;
        if (bcast) s = __ldcg(mine + (size_t)pp.root * kPeerSlotDoubles + i);
        else {
            s = __ldcg(mine + i);
            for (int r = 1; r < pv.world; r++) s += __ldcg(mine + (size_t)r * kPeerSlotDoubles + i);
        }
        pp.buf[i] = s;
    }
}

}  // namespace

// Enough CTAs to give every right-hand side its own and every warp only a few elements to reduce; never more than the
// SMs (the CTAs spin on each other: they must all be resident).
int tail_blocks(int n_elem, int G, int num_sms) {
    const int want = std::max((G + 3) / 4, (n_elem + 15) / 16);   // four warps per CTA: four right-hand sides, sixteen elements
    return std::max(1, std::min(want, num_sms));
}

void launch_tail(const TailParams& tp, int blocks, cudaStream_t stream) {
    const int K = tp.solve.K;
    const size_t smem = tp.do_solve ? solve_smem_bytes(K) : 0;
    if (K <= 32) {
        if (ensure_dynamic_smem((const void*)tail_kernel<1>, smem) != cudaSuccess) return;
        launch_pdl(tp.pdl != 0, tail_kernel<1>, blocks, kTailThreads, smem, stream, tp);
    } else {
        if (ensure_dynamic_smem((const void*)tail_kernel<2>, smem) != cudaSuccess) return;
        launch_pdl(tp.pdl != 0, tail_kernel<2>, blocks, kTailThreads, smem, stream, tp);
    }
}

void launch_peer_small(const PeerSmallParams& pp, cudaStream_t stream) { peer_small_kernel<<<1, 256, 0, stream>>>(pp); }

}  // namespace lsmc
