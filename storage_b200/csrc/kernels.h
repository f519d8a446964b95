// kernels.h — launchers of the sm_100a kernels (kernels.cu).
#pragma once
#include <algorithm>
#include <cstddef>
#include <utility>
#include "../../include/lsmc_b200.h"
#include "lsmc_device.cuh"

struct CUtensorMap_st;   // <cuda.h>: CUtensorMap

namespace lsmc {

struct SweepConfig {
    int TP;        // paths per tile (56, 40 or 24)
    int MT;        // basis m-tiles: 1, 2, 4, 7 or 8 (8*MT >= K)
    int blocks;    // persistent CTAs
    int n_scratch; // rows of basis-program nodes that are not basis functions (host_plan.h RowProgram)
    int nbuf;      // shared-memory buffers for the staged V_next tile (2 = prefetch the next tile)
    int n_rowops;  // multiplies of the row program
};

struct SolveParams {
    const double* reduced;   // [8*MT][ncols]
    int ncols, K, G, NT, ldb, accum_mode;
    int gram_lower_blocks;   // 1: of X~^T X~ only the blocks (rows 8m.., columns 8kb..) with m <= kb are filled (general sweep)
    int64_t n_global;
    double pivot_tol;
    const double* precond;   // T [K][K] + flags [K] of the regressed period
    double* beta;            // [G][ldb]
};

struct FinalParams {
    int64_t N;
    int pass, must_be_empty, paired;
    double tc0, tc1, tc2, tc3, mean;
    const double* spot_end;
    const double* inventory;
    double* pv_by_sim;
    double* sums;            // [gridDim.x][3]
    double* o_inventory; double* o_pv;
};

// Opt a kernel in to `bytes` of dynamic shared memory on the CURRENT device.  The attribute is per (kernel, device):
// the sizes already granted are cached under a mutex, so contexts on several GPUs (or several host threads) each get it.
cudaError_t ensure_dynamic_smem(const void* kernel, size_t bytes);

// Launch with programmatic stream serialisation (lsmc_device.cuh pdl_wait / pdl_launch_dependents).  LSMC_PDL=0 in the
// environment falls back to plain stream order (development).
bool pdl_enabled();
// TMA descriptor of an FP64 panel [rows][Npad]: box = TP paths x box_rows rows (sweep4.cu).  False when the driver lacks the encoder.
bool make_panel_map(CUtensorMap_st* map, const double* base, int64_t Npad, int rows, int TP, int box_rows);
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(bool pdl, void (*kern)(KArgs...), int blocks, int threads, size_t smem, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)blocks); cfg.blockDim = dim3((unsigned)threads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}

bool choose_sweep_shape(int max_grid, int K, int F, int MT, int n_scratch, int n_rowops, int* TP, int* nbuf);
int sweep_max_items();
int precond_rows(int K, int64_t N);
void launch_precond(const PrecondParams& q, cudaStream_t stream);
void launch_simulate(const SimParams& sp, cudaStream_t stream);
cudaError_t launch_sweep(SweepParams& q, const SweepConfig& cfg, cudaStream_t stream);
// pipelined sweep of the regular periods (sweep4.cu)
int sweep4_tile(int max_grid, int K, int F, int n_prod);
bool sweep4_covers(const SweepParams& q);
cudaError_t launch_sweep4(SweepParams& q, int TP, int n_prod, int blocks, cudaStream_t stream);
void launch_reduce_partials(const double* partial, int n_parts, int n_elem, double* out, cudaStream_t stream);
void launch_solve(const SolveParams& sp, cudaStream_t stream);
void launch_forward(const ForwardParams* device_params, int p0, int p1, int K, int n_nodes, int max_grid, int blocks, bool fast, cudaStream_t stream);
size_t forward_smem_bytes(int K, int n_nodes, int max_grid, bool fast);
int forward_fast_blocks_per_sm();
int forward_fast_threads();
// Streaming forward kernel (forward2.cu): per-period constants addressed by the period index alone.
constexpr int kFwd2PillarDoubles = 80;   // per period: pillar inventories x[16] (+inf padded), min rates, max rates, slopes of both
__host__ __device__ constexpr int forward2_beta_pitch(int K) { return (((K + 1) & ~1) / 2) % 2 ? ((K + 1) & ~1) : ((K + 1) & ~1) + 2; }   // even, odd half: conflict-free 128-bit rows
struct Forward2Args {
    const ForwardParams* params;   // [P-1]
    int p0, p1;                    // periods of this launch
    int K, F, n_nodes, max_grid, ldb, grid_pitch, n_parts;
    const BasisDev* basis;
    const double* beta_all; size_t beta_stride;   // coefficients of period p at beta_all + p * beta_stride, rows of pitch ldb
    const double* grid_pad;        // [P][grid_pitch] the grid of period p at row p
    const double* pillars;         // [P-1][kFwd2PillarDoubles]
    double* warp_sums;             // [P-1][n_parts][kFwdSums + K] one partial row per warp
};
cudaError_t launch_forward2(const Forward2Args& a, int K, int blocks, cudaStream_t stream);
size_t forward2_smem_bytes(int K, int n_nodes, int max_grid);
int forward2_blocks_per_sm(int K);
int forward2_threads();
int forward2_parts(int blocks);
void launch_reduce_rows(const double* in, int n_rows, int n_blocks, int n_cols, double* out, cudaStream_t stream);
void launch_final(const FinalParams& q, int blocks, cudaStream_t stream);
void launch_fill(double* p, int64_t n, double v, cudaStream_t stream);

}  // namespace lsmc
