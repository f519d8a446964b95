// lsmc_api.cu — the C ABI (include/lsmc_b200.h): context, device memory, per-period orchestration.
// Replaces the body of LsmcStorageValuation.Calculate<T> (LSV.cs:97-616) behind plain pointers and sizes.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <limits>
#include <string>
#include <vector>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>

#include "host_plan.h"
#include "kernels.h"
#include "mt19937.h"
#include "peer.h"

using namespace lsmc;

namespace {

thread_local std::string g_create_error;

#define CUDA_CHECK(expr)                                                                              \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess)                                                                        \
            throw Error(LSMC_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));              \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    void alloc(size_t count) {
        if (count <= n && p) return;
        release();
        CUDA_CHECK(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
        n = count;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    void upload(const std::vector<T>& h, cudaStream_t s) {
        alloc(h.size());
        if (!h.empty()) CUDA_CHECK(cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s));
    }
    ~DevBuf() { release(); }
};

// Page-locked host staging for the small read-backs of a valuation.  A cudaMemcpyAsync to PAGEABLE memory blocks inside
// the runtime until the stream has drained; with several ranks in one process (one host thread per GPU) the stream may be
// waiting for a peer whose host thread then cannot launch — every device-to-host copy between the first and the last
// exchange of a valuation therefore lands here and is followed by a plain cudaStreamSynchronize.
struct PinnedBuf {
    double* p = nullptr;
    size_t n = 0;
    void alloc(size_t count) {
        if (count <= n && p) return;
        release();
        CUDA_CHECK(cudaMallocHost(&p, count * sizeof(double)));
        n = count;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; n = 0; }
    ~PinnedBuf() { release(); }
};
constexpr size_t kPinScalars = 16;   // [0] backward sum, [1..3] final sums, [4..6] squared deviations, [8] error flag (as int)

// ---- NCCL through dlopen: the library has no link-time dependency on it --------------------------------------
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, char[128], int) = nullptr;   // ncclUniqueId is a 128-byte struct passed by value
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool load(std::string& err) {
        if (handle) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) { handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (handle) break; }
        if (!handle) { err = "cannot dlopen libnccl.so.2"; return false; }
        GetUniqueId = (int (*)(void*))dlsym(handle, "ncclGetUniqueId");
        CommInitRank = (int (*)(void**, int, char[128], int))dlsym(handle, "ncclCommInitRank");
        AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(handle, "ncclAllReduce");
        Broadcast = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(handle, "ncclBroadcast");
        AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(handle, "ncclAllGather");
        CommDestroy = (int (*)(void*))dlsym(handle, "ncclCommDestroy");
        GetErrorString = (const char* (*)(int))dlsym(handle, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !AllReduce || !Broadcast || !AllGather || !CommDestroy) { err = "libnccl is missing symbols"; return false; }
        return true;
    }
};
NcclApi g_nccl;
struct NcclId { char bytes[128]; };

// The exchange buffer of one rank and its view of the other ranks' buffers (peer.h).
struct PeerComm {
    bool ready = false;               // every rank's buffer is mapped here
    double* buf = nullptr;            // this rank's allocation: slots [2][world][kPeerSlotDoubles] + the arrival counter
    int alloc_world = 0;
    PeerView view{};
    unsigned long long target = 0;    // cumulative arrivals expected at this rank's counter
    unsigned long long local_target = 0;   // cumulative tickets of this rank's own CTAs
    unsigned long long epoch = 0;     // exchanges started (the same sequence on every rank)
    std::vector<void*> opened;        // cudaIpcOpenMemHandle mappings to close

    static size_t bytes(int world) { return (size_t)2 * world * kPeerSlotDoubles * sizeof(double) + 256; }
    unsigned long long* counter_of(double* base, int world) const { return reinterpret_cast<unsigned long long*>(base + (size_t)2 * world * kPeerSlotDoubles); }
    unsigned long long* ticket_of(double* base, int world) const { return counter_of(base, world) + 16; }   // own GPU only, its own 128-byte line
    void allocate(int world) {
        if (buf && alloc_world == world) return;
        release();
        CUDA_CHECK(cudaMalloc(&buf, bytes(world)));
        CUDA_CHECK(cudaMemset(buf, 0, bytes(world)));
        alloc_world = world;
        target = 0; local_target = 0; epoch = 0;
    }
    void self_only() {   // one rank: the "exchange" is the grid barrier of the tail kernel
        allocate(1);
        view = PeerView{};
        view.rank = 0; view.world = 1; view.slots[0] = buf; view.counter[0] = counter_of(buf, 1); view.ticket = ticket_of(buf, 1);
        ready = true;
    }
    void release() {
        for (void* p : opened) cudaIpcCloseMemHandle(p);
        opened.clear();
        if (buf) cudaFree(buf);
        buf = nullptr; alloc_world = 0; ready = false;
    }
    // parity and cumulative target of the next exchange in which every rank arrives with `ctas` CTAs
    void next(int ctas, int* parity, unsigned long long* tgt, unsigned long long* local_tgt) {
        *parity = (int)(epoch & 1);
        epoch++;
        target += (unsigned long long)ctas * (unsigned long long)view.world;
        local_target += (unsigned long long)ctas;
        *tgt = target; *local_tgt = local_target;
    }
    ~PeerComm() { release(); }
};

// Ranks that live in one process (one host thread each): shared cancellation flag and a host barrier.
struct LocalGroup {
    std::atomic<int> cancelled{0};
    std::mutex mu;
    std::condition_variable cv;
    int waiting = 0, generation = 0;
    void barrier(int n) {
        std::unique_lock<std::mutex> lock(mu);
        const int gen = generation;
        if (++waiting == n) { waiting = 0; generation++; cv.notify_all(); }
        else cv.wait(lock, [&] { return generation != gen; });
    }
};

}  // namespace

struct lsmc_ctx {
    int device = 0, num_sms = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    bool has_problem = false, sims_ready[2] = {false, false}, backward_done = false, forward_done = false;
    int fac_set = -1;   // which simulation set currently owns the shared factor panel `fac` (-1: none)
    Plan plan;
    int64_t N = 0, Npad = 0, Nglobal = 0;
    int TP = 56, nbuf = 2, MT = 1, ldb = 4, sweep_blocks = 0, fwd_blocks = 0, ncols_max = 0;
    bool fwd_fast = false;   // straight-line forward kernel (forward_fast_ok)
    bool fwd2 = false;       // its streaming form (forward2.cu): additionally every ratchet table has at most 16 pillars
    int fwd_parts = 0;       // partial rows of the forward sums per period (CTAs, or warps with forward2)
    int TP4 = 0, sweep4_blocks = 0;   // pipelined sweep (sweep4.cu): tile size (0 = shape not covered) and persistent CTAs
    uint32_t outputs = 0;       // bit0 per-sim panels, bit1 decision index
    bool keep_values = false;
    int valuation_antithetic = 0;

    DevBuf<double> spot[2], fac, Vbuf, Vall, beta_all, grid, partial, reduced, inventory, pv_by_sim, fwd_sums, fwd_red, fin_sums;
    DevBuf<double> t_invcost, t_vol, t_cost, t_cons, t_wlo, t_whi, c_x, c_lo, c_hi;
    DevBuf<int32_t> t_ndec, t_lo, t_hi;
    DevBuf<double> s_decay, s_chol, s_vol, s_drift, s_fwd, normals_stage;
    DevBuf<BasisDev> basis;
    DevBuf<ForwardParams> fwd_params;   // [P-1] per-period descriptors of the forward pass
    DevBuf<double> grid_pad, fwd_pillars;   // forward2: [P][grid_pitch] grids, [P-1][kFwd2PillarDoubles] ratchet pillars + slopes
    DevBuf<double> precond;     // [S][precond_stride(K)] regression preconditioners (rank 0's, broadcast)
    DevBuf<int> error_flag;
    DevBuf<double> panels[6];
    DevBuf<int8_t> decisions;
    std::vector<char> beta_override_set;   // [P-1]
    std::vector<double> beta_override;     // [P-1][max_grid][ldb]

    // multi-GPU
    void* comm = nullptr;
    int rank = 0, world = 1;
    PeerComm peer;                       // exchange over peer memory (world == 1: the tail kernel's grid barrier)
    Mt19937 mt;                          // LSMC_RNG_MT19937: the reference's seeded generator, kept so that the valuation set can continue it
    std::vector<double> mt_normals;      // its normals for this rank's paths, [S][F][N]
    bool use_pdl = true;                 // programmatic dependent launch of the backward loop's kernels.  Off when another rank of the
                                         // group runs on the same GPU: its CTAs, parked at their dependency wait, could hold the SMs that
                                         // the other rank's kernels — which this rank's exchange is waiting for — need to become resident.
    bool time_kernels = false;           // lsmc_set_kernel_timing: CUDA events around every sweep launch (they break the kernels' programmatic overlap)
    bool use_tail = true;                // fused reduce + exchange + solve kernel (tail.cu); false: reduce -> NCCL -> solve
    std::shared_ptr<LocalGroup> group;   // ranks of one process (lsmc_comm_init_local)
    int64_t out_pitch = 0, out_col0 = 0; // per-sim outputs of lsmc_get_results: row pitch / first column in the caller's buffers (0: n_sims, 0)

    // results staged on host
    double backward_npv = 0.0, npv = 0.0, npv_se = 0.0, final_inventory_sum = 0.0, terminal_pv_sum = 0.0;
    std::vector<double> h_fwd;   // [P-1][nsum]
    PinnedBuf pin;               // kPinScalars scalars, then the [P-1][nsum] forward sums
    PinnedBuf export_stage;      // two page-locked staging chunks of the panel export (pageable destinations)
    bool export_pending = false; // asynchronous copies into page-locked destinations not yet synchronised
    cudaEvent_t export_ev[2] = {nullptr, nullptr};

    // timings
    lsmc_timings tm{};
    std::vector<cudaEvent_t> ev;

    ~lsmc_ctx() {
        for (auto e : ev) cudaEventDestroy(e);
        for (auto e : export_ev) if (e) cudaEventDestroy(e);
        if (comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm);
        if (stream) cudaStreamDestroy(stream);
    }
    size_t beta_stride() const { return (size_t)(plan.max_grid + 8) * ldb; }
    double* V(int which) { return Vbuf.p + (size_t)which * plan.max_grid * Npad; }
    cudaEvent_t event(size_t i) {
        while (ev.size() <= i) { cudaEvent_t e; CUDA_CHECK(cudaEventCreate(&e)); ev.push_back(e); }
        return ev[i];
    }
    void peer_small(double* buf, size_t count, int root) {
        for (size_t o = 0; o < count; o += kPeerSlotDoubles) {
            PeerSmallParams pp{};
            pp.buf = buf + o; pp.count = (int)std::min<size_t>(kPeerSlotDoubles, count - o); pp.root = root;
            pp.peers = peer.view; pp.error_flag = error_flag.p;
            peer.next(1, &pp.parity, &pp.target, &pp.local_target);
            launch_peer_small(pp, stream);
            tm.kernel_launches++;
        }
    }
    void allreduce(double* buf, size_t count) {
        if (world <= 1) return;
        if (peer.ready && (use_tail || !comm)) { peer_small(buf, count, -1); return; }
        int rc = g_nccl.AllReduce(buf, buf, count, /*ncclFloat64*/ 8, /*ncclSum*/ 0, comm, stream);
        if (rc != 0) throw Error(LSMC_ECUDA, std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"));
    }
    void broadcast0(double* buf, size_t count) {
        if (world <= 1) return;
        if (peer.ready && (use_tail || !comm)) { peer_small(buf, count, 0); return; }
        int rc = g_nccl.Broadcast(buf, buf, count, /*ncclFloat64*/ 8, /*root*/ 0, comm, stream);
        if (rc != 0) throw Error(LSMC_ECUDA, std::string("ncclBroadcast: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"));
    }
};

namespace {

// development (LSMC_TRACE_HOST): host-side checkpoints of every rank on stderr
static const bool g_trace_host = getenv("LSMC_TRACE_HOST") != nullptr;
#define HOST_TRACE(c, what)                                                                                                       \
    do {                                                                                                                          \
        if (g_trace_host)                                                                                                         \
            fprintf(stderr, "[rank %d %.6f] %s\n", (c)->rank,                                                                     \
                    std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(), what);            \
    } while (0)

template <typename Fn>
int guarded(lsmc_ctx* ctx, Fn&& fn) {
    try {
        if (ctx) CUDA_CHECK(cudaSetDevice(ctx->device));
        fn();
        return LSMC_OK;
    } catch (const Error& e) {
        if (ctx) ctx->err = e.what(); else g_create_error = e.what();
        return e.status;
    } catch (const std::exception& e) {
        if (ctx) ctx->err = e.what(); else g_create_error = e.what();
        return LSMC_EINVAL;
    }
}

void require(bool ok, int status, const char* msg) { if (!ok) throw Error(status, msg); }

void check_device_error_flag(lsmc_ctx* c) {
    int* h_flag = reinterpret_cast<int*>(c->pin.p + 8);
    CUDA_CHECK(cudaMemcpyAsync(h_flag, c->error_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUDA_CHECK(cudaStreamSynchronize(c->stream));
    const int flag = *h_flag;
    switch (flag) {
        case 0: return;
        case 1: case 2: throw Error(LSMC_EINVAL, "A simulated inventory is outside the storage's allowed inventory range (CmdtyStorage.GetInjectWithdrawRange).");
        case 3: throw Error(LSMC_EINVAL, "Inventory constraints cannot be fulfilled. This could potentially be fixed by increasing the numerical tolerance.");
        case kPeerTimeout: {
            unsigned long long have = 0;
            cudaMemcpy(&have, c->peer.view.counter[c->peer.view.rank], sizeof(have), cudaMemcpyDeviceToHost);
            throw Error(LSMC_ECUDA, "The exchange between the GPUs timed out: another rank failed, was cancelled or is not running the same valuation (rank " +
                        std::to_string(c->rank) + " of " + std::to_string(c->world) + ", exchange " + std::to_string(c->peer.epoch) + ", arrivals " +
                        std::to_string(have) + " of " + std::to_string(c->peer.target) + ").");
        }
        default: throw Error(LSMC_EINVAL, "Inventory is outside of inventoryGrid bounds.");
    }
}

// One process per GPU: every rank exports its exchange buffer with cudaIpcGetMemHandle, the handles travel through one
// ncclAllGather, every rank maps the others' buffers.  All ranks then agree (all-reduce of a flag) on whether the peer path
// is usable; if any rank could not map a peer, all of them keep the NCCL collectives.
void map_peers_over_ipc(lsmc_ctx* c) {
    const int world = c->world, rank = c->rank;
    c->peer.release();
    if (world > kMaxPeers) return;
    if (const char* e = getenv("LSMC_COMM")) { if (std::string(e) == "nccl") return; }   // development: NCCL collectives only
    cudaStream_t s = c->stream;
    bool ok = true;
    try { c->peer.allocate(world); } catch (const Error&) { ok = false; }
    struct Blob { cudaIpcMemHandle_t h; int ok; int pad[3]; };
    static_assert(sizeof(Blob) % 16 == 0, "blob size");
    DevBuf<unsigned char> d_send, d_recv;
    d_send.alloc(sizeof(Blob)); d_recv.alloc(sizeof(Blob) * world);
    Blob mine{};
    if (ok && cudaIpcGetMemHandle(&mine.h, c->peer.buf) != cudaSuccess) { (void)cudaGetLastError(); ok = false; }
    mine.ok = ok ? 1 : 0;
    {   // which GPU this rank runs on: two ranks of one GPU must not use programmatic dependent launch (lsmc_ctx::use_pdl)
        int dom = 0, bus = 0, dev = 0;
        cudaDeviceGetAttribute(&dom, cudaDevAttrPciDomainId, c->device);
        cudaDeviceGetAttribute(&bus, cudaDevAttrPciBusId, c->device);
        cudaDeviceGetAttribute(&dev, cudaDevAttrPciDeviceId, c->device);
        mine.pad[0] = (dom << 16) | (bus << 8) | dev;
    }
    CUDA_CHECK(cudaMemcpyAsync(d_send.p, &mine, sizeof(Blob), cudaMemcpyHostToDevice, s));
    int rc = g_nccl.AllGather(d_send.p, d_recv.p, sizeof(Blob), /*ncclInt8*/ 0, c->comm, s);
    if (rc != 0) throw Error(LSMC_ECUDA, "ncclAllGather failed while exchanging the peer-memory handles");
    std::vector<Blob> all(world);
    CUDA_CHECK(cudaMemcpyAsync(all.data(), d_recv.p, sizeof(Blob) * world, cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    PeerView v{};
    v.rank = rank; v.world = world;
    for (int r = 0; r < world; r++) if (r != rank && all[r].pad[0] == mine.pad[0]) c->use_pdl = false;
    for (int r = 0; r < world && ok; r++) {
        if (!all[r].ok) { ok = false; break; }
        double* base = c->peer.buf;
        if (r != rank) {
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { (void)cudaGetLastError(); ok = false; break; }
            c->peer.opened.push_back(p);
            base = static_cast<double*>(p);
        }
        v.slots[r] = base;
        v.counter[r] = c->peer.counter_of(base, world);
    }
    v.ticket = c->peer.ticket_of(c->peer.buf, world);
    // agreement: the minimum of the ranks' flags
    DevBuf<double> flag;
    flag.alloc(1);
    const double f = ok ? 1.0 : 0.0;
    CUDA_CHECK(cudaMemcpyAsync(flag.p, &f, sizeof(double), cudaMemcpyHostToDevice, s));
    rc = g_nccl.AllReduce(flag.p, flag.p, 1, /*ncclFloat64*/ 8, /*ncclMin*/ 3, c->comm, s);
    if (rc != 0) throw Error(LSMC_ECUDA, "ncclAllReduce failed while agreeing on the peer-memory path");
    double agreed = 0.0;
    CUDA_CHECK(cudaMemcpyAsync(&agreed, flag.p, sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    if (agreed == 1.0) { c->peer.view = v; c->peer.ready = true; }
    else c->peer.release();
    if (getenv("LSMC_VERBOSE")) fprintf(stderr, "lsmc_comm_init: rank %d of %d, exchange over %s\n", rank, world, c->peer.ready ? "peer memory (cudaIpc)" : "NCCL collectives");
}

// The straight-line forward kernels (forward2.cu; kernels.cu forward_fast_kernel for K <= 16) cover: no extra decisions and next-period
// grids on which its O(1) lookup provably equals the bisection of SH.cs:280-314 (kernels.cu fwd_class_result): at most 127
// points; every interval but the last wider than twice the tolerance (the last one only has to be positive: accumulating
// the step and clamping to the maximum, FixedSpacingStateSpaceGridCalc.cs:45-62, can leave one of a rounding error); all
// points but the last within a quarter step of fixed spacing, so that the index guess is off by at most one.
bool forward_fast_ok(const Plan& pl, const lsmc_problem& pb) {
    if (pb.extra_decisions != 0 || pl.K > 64 || pl.max_dec > 3) return false;
    const double tol = pb.numerical_tolerance;
    if (!(tol >= 0.0)) return false;
    for (int p = 1; p < pl.P; p++) {
        const int n = pl.grid_size(p);
        const double* g = pl.grid_of(p);
        if (n > 127) return false;
        if (n < 2) continue;
        const double h = g[1] - g[0];
        if (!(h > 0.0) || !std::isfinite(1.0 / h)) return false;
        if (!(g[n - 1] > g[n - 2]) || !(g[n - 1] - g[n - 2] <= 1.25 * h)) return false;
        for (int l = 0; l + 1 < n; l++) {
            if (l + 2 < n && !(g[l + 1] - g[l] > 2.0 * tol)) return false;
            if (!(std::fabs(g[l] - (g[0] + l * h)) <= 0.25 * h)) return false;
        }
    }
    return true;
}

// ---------------------------------------------------------------------------------------------------------------
void set_problem(lsmc_ctx* c, const lsmc_problem* problem) {
    require(problem != nullptr, LSMC_EINVAL, "problem is null");
    c->has_problem = false;
    const auto tp0 = std::chrono::steady_clock::now();
    build_plan(*problem, c->plan);
    if (getenv("LSMC_VERBOSE")) fprintf(stderr, "set_problem: build_plan %.2f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp0).count());
    Plan& pl = c->plan;
    const lsmc_problem& pb = pl.problem.pb;
    c->N = pb.n_sims;
    c->Nglobal = pb.n_sims_global > 0 ? pb.n_sims_global : pb.n_sims;
    if (c->Nglobal != c->N)   // a shard of a larger valuation: antithetic partners 2m, 2m+1 must live on one rank
        require(pb.sim_offset % 2 == 0 && pb.sim_offset >= 0 && pb.sim_offset + c->N <= c->Nglobal && (c->N % 2 == 0 || pb.sim_offset + c->N == c->Nglobal),
                LSMC_EINVAL, "a shard must start on an even path and hold an even number of paths (except the last shard)");
    c->MT = pl.K <= 8 ? 1 : (pl.K <= 16 ? 2 : (pl.K <= 32 ? 4 : (pl.K > 48 && pl.K <= 56 ? 7 : 8)));   // sweep.cu instantiates 1, 2, 4, 7, 8
    c->ldb = (pl.K + 3) & ~3;
    require(pl.basis.n_nodes <= kMaxNodes && (int)pl.basis.ops.size() <= kMaxOps, LSMC_EUNSUPPORTED, "basis functions need too many product nodes");
    require(pl.max_dec <= 16, LSMC_EUNSUPPORTED, "extra_decisions above 6 is not supported on the device path");
    require(choose_sweep_shape(pl.max_grid, pl.K, pl.F, c->MT, pl.basis_rows.n_scratch, (int)pl.basis_rows.ops.size(), &c->TP, &c->nbuf), LSMC_EUNSUPPORTED,
            "inventory grid / basis too large for the sweep kernel's shared memory");
    require((pl.max_grid + 7) / 8 + c->MT <= sweep_max_items(), LSMC_EUNSUPPORTED, "inventory grid has too many points for the sweep kernel (about 200 at most)");
    const int64_t n_tiles = (c->N + c->TP - 1) / c->TP;
    c->Npad = (c->N + 3359) / 3360 * 3360;   // a multiple of every sweep tile size (lcm of 56, 48, 40, 32, 24)
    c->sweep_blocks = (int)std::min<int64_t>(c->num_sms, n_tiles);
    c->TP4 = sweep4_tile(pl.max_grid, pl.K, pl.F, (int)pl.basis.ops.size());
    c->sweep4_blocks = c->TP4 ? (int)std::min<int64_t>(c->num_sms, (c->N + c->TP4 - 1) / c->TP4) : 0;
    if (const char* e = getenv("LSMC_SWEEP")) { if (atoi(e) == 3) c->TP4 = c->sweep4_blocks = 0; }   // development: force the general sweep kernel
    if (getenv("LSMC_TIME_KERNELS")) c->time_kernels = true;   // development: per-launch events without the API call
    if (const char* e = getenv("LSMC_TAIL")) c->use_tail = atoi(e) != 0;   // development: 0 = reduce -> NCCL all-reduce -> one-CTA solve
    if (c->world <= 1 && !c->peer.ready) c->peer.self_only();
    require(8 * c->MT * 8 * ((pl.max_grid + 7) / 8 + c->MT) <= kPeerSlotDoubles, LSMC_EUNSUPPORTED, "normal-equation block larger than the exchange slots");
    bool fwd_grid_ok = forward_fast_ok(pl, pb);
    if (const char* e = getenv("LSMC_FORWARD")) { if (atoi(e) == 0) fwd_grid_ok = false; }   // development: force the general forward kernel
    c->fwd_fast = fwd_grid_ok && pl.K <= 16;
    require(forward_smem_bytes(pl.K, pl.basis.n_nodes, pl.max_grid, c->fwd_fast) <= (size_t)227 * 1024 - 4096, LSMC_EUNSUPPORTED,
            "basis functions / inventory grid too large for the forward kernel's shared memory");
    c->fwd2 = fwd_grid_ok;   // the streaming kernel also has a K <= 64 form (forward2.cu KT = 0)
    for (int cid = 0; cid < pb.n_constraints; cid++) c->fwd2 = c->fwd2 && pb.constraint_count[cid] <= 16;
    c->fwd2 = c->fwd2 && forward2_smem_bytes(pl.K, pl.basis.n_nodes, pl.max_grid) <= (pl.K <= 16 ? (size_t)75 * 1024 : (size_t)227 * 1024 - 2048);
    if (const char* e = getenv("LSMC_FORWARD")) { if (atoi(e) == 1) c->fwd2 = false; }   // development: 1 = the previous straight-line kernel
    const int fwd_threads = c->fwd2 ? forward2_threads() : (c->fwd_fast ? forward_fast_threads() : 256);
    const int fwd_per_sm = c->fwd2 ? forward2_blocks_per_sm(pl.K) : (c->fwd_fast ? forward_fast_blocks_per_sm() : (pl.K <= 16 ? 3 : 1));
    c->fwd_blocks = (int)std::min<int64_t>((int64_t)c->num_sms * fwd_per_sm, (c->N + fwd_threads - 1) / fwd_threads);   // all CTAs resident: each walks every period
    c->fwd_parts = c->fwd2 ? forward2_parts(c->fwd_blocks) : c->fwd_blocks;
    c->ncols_max = 8 * ((pl.max_grid + 7) / 8 + c->MT);

    cudaStream_t s = c->stream;
    const size_t S = pl.S, F = pl.F, Np = (size_t)c->Npad;
    // Only the pad columns [N, Npad) of every row have to be finite zeros (the last tile of a sweep reads them, and a
    // NaN there would poison the normal equations); columns [0, N) are always written before they are read.
    auto zero_pad = [&](double* base, size_t rows) {
        const size_t pad = Np - (size_t)c->N;
        if (pad && rows) CUDA_CHECK(cudaMemset2DAsync(base + c->N, Np * sizeof(double), 0, pad * sizeof(double), rows, s));
    };
    for (int set = 0; set < 2; set++) { c->spot[set].alloc(S * Np); zero_pad(c->spot[set].p, S); }
    c->fac.alloc(S * F * Np); zero_pad(c->fac.p, S * F);
    c->Vbuf.alloc((size_t)2 * pl.max_grid * Np);
    zero_pad(c->Vbuf.p, (size_t)2 * pl.max_grid);
    if (c->keep_values) { c->Vall.alloc((size_t)pl.P * pl.max_grid * Np); zero_pad(c->Vall.p, (size_t)pl.P * pl.max_grid); }
    c->beta_all.alloc((size_t)(pl.P - 1) * c->beta_stride());
    CUDA_CHECK(cudaMemsetAsync(c->beta_all.p, 0, c->beta_all.n * sizeof(double), s));
    c->grid.upload(pl.grid, s);
    c->partial.alloc((size_t)std::max(c->sweep_blocks, c->sweep4_blocks) * 8 * c->MT * c->ncols_max);
    CUDA_CHECK(cudaMemsetAsync(c->partial.p, 0, c->partial.n * sizeof(double), s));
    c->reduced.alloc((size_t)8 * c->MT * c->ncols_max);
    c->inventory.alloc(Np); c->pv_by_sim.alloc(Np);
    const int nsum = kFwdSums + pl.K;
    c->fwd_sums.alloc((size_t)(pl.P - 1) * c->fwd_parts * nsum);
    if (c->fwd2) {   // per-period constants of the streaming forward kernel, addressed by the period index alone
        const int gp = (pl.max_grid + 1) & ~1;
        std::vector<double> hg((size_t)pl.P * gp, 0.0), hp((size_t)(pl.P - 1) * kFwd2PillarDoubles, 0.0);
        const double inf = std::numeric_limits<double>::infinity();
        for (int p = 0; p < pl.P; p++) std::copy(pl.grid_of(p), pl.grid_of(p) + pl.grid_size(p), hg.begin() + (size_t)p * gp);
        for (int p = 0; p < pl.P - 1; p++) {
            double* x = &hp[(size_t)p * kFwd2PillarDoubles];
            double *lo = x + 16, *hi = x + 32, *slo = x + 48, *shi = x + 64;
            const int cid = pb.constraint_id[p], first = pb.constraint_first[cid], n = pb.constraint_count[cid], kind = pb.constraint_kind[cid];
            const double* px = pb.pillar_inventory + first; const double* plo = pb.pillar_min_rate + first; const double* phi = pb.pillar_max_rate + first;
            for (int i = 0; i < 16; i++) {
                // piecewise linear: interior pillars only take part in the segment count (entries from n - 1 on are +inf);
                // step: every pillar does (entries from n on are +inf)
                const int last_real = kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR ? n - 1 : n;
                x[i] = i < last_real ? px[i] : inf;
                lo[i] = i < n ? plo[i] : 0.0; hi[i] = i < n ? phi[i] : 0.0;
                if (kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR && i + 1 < n) {   // slopes as LinearSpline forms them (kernels.cu spline_dev)
                    const double dx = px[i + 1] - px[i];
                    slo[i] = (plo[i + 1] - plo[i]) / dx; shi[i] = (phi[i + 1] - phi[i]) / dx;
                }
            }
            if (kind == LSMC_CONSTRAINT_CONSTANT) x[0] = 0.0;
        }
        c->grid_pad.upload(hg, s); c->fwd_pillars.upload(hp, s);
        CUDA_CHECK(cudaStreamSynchronize(s));   // hg, hp go out of scope
    }
    c->fwd_red.alloc((size_t)(pl.P - 1) * nsum);
    c->fin_sums.alloc((size_t)c->fwd_blocks * 3);
    c->t_ndec.upload(pl.tables.n_dec, s); c->t_invcost.upload(pl.tables.inv_cost_npv, s);
    c->t_vol.upload(pl.tables.vol, s); c->t_cost.upload(pl.tables.cost_npv, s); c->t_cons.upload(pl.tables.consumed, s);
    c->t_wlo.upload(pl.tables.w_lo, s); c->t_whi.upload(pl.tables.w_hi, s);
    c->t_lo.upload(pl.tables.lo, s); c->t_hi.upload(pl.tables.hi, s);
    c->c_x.upload(pl.problem.pillar_inventory, s); c->c_lo.upload(pl.problem.pillar_min_rate, s); c->c_hi.upload(pl.problem.pillar_max_rate, s);
    c->s_decay.upload(pl.sim.decay, s); c->s_chol.upload(pl.sim.chol, s); c->s_vol.upload(pl.sim.vol, s);
    c->s_drift.upload(pl.sim.drift, s); c->s_fwd.upload(pl.sim.fwd, s);
    {
        std::vector<BasisDev> hb(1);
        BasisDev& b = hb[0];
        std::memset(&b, 0, sizeof(b));
        b.n_nodes = pl.basis.n_nodes; b.n_ops = (int)pl.basis.ops.size(); b.K = pl.K; b.F = pl.F;
        for (size_t i = 0; i < pl.basis.ops.size(); i++) b.ops[i] = BasisOpDev{pl.basis.ops[i].dst, pl.basis.ops[i].a, pl.basis.ops[i].b, 0};
        for (int k = 0; k < pl.K; k++) {
            b.node_of_basis[k] = pl.basis.node_of_basis[k];
            require(pb.spot_power[k] <= 127, LSMC_EUNSUPPORTED, "basis exponent above 127");
            b.spot_pow[k] = (int8_t)pb.spot_power[k];
            for (int f = 0; f < pl.F; f++) {
                require(pb.factor_power[k * pl.F + f] <= 127, LSMC_EUNSUPPORTED, "basis exponent above 127");
                b.fac_pow[k][f] = (int8_t)pb.factor_power[k * pl.F + f];
            }
        }
        {   // waves of the product program (sweep4.cu): an op joins the current wave unless it reads a result of that wave
            int nw = 0, w0 = 0;
            for (int o = 0; o < b.n_ops; o++) {
                bool dep = false;
                for (int e = w0; e < o; e++) dep = dep || b.ops[e].dst == b.ops[o].a || b.ops[e].dst == b.ops[o].b;
                if (dep) { b.opwave_end[nw++] = (int16_t)o; w0 = o; }
            }
            if (b.n_ops > 0) b.opwave_end[nw++] = (int16_t)b.n_ops;
            b.n_opwaves = nw;
        }
        const RowProgram& rp = pl.basis_rows;
        require((int)rp.ops.size() <= kMaxRowOps, LSMC_EUNSUPPORTED, "basis functions need too many product nodes");
        b.n_rowops = (int)rp.ops.size(); b.n_rowwaves = (int)rp.wave_end.size(); b.n_scratch = rp.n_scratch;
        for (size_t i = 0; i < rp.wave_end.size(); i++) b.rowwave_end[i] = rp.wave_end[i];
        for (size_t i = 0; i < rp.ops.size(); i++)
            b.rowops[i] = RowOpDev{rp.ops[i].ra, rp.ops[i].rb, rp.ops[i].rd, 0, rp.ops[i].a, rp.ops[i].b, rp.ops[i].d, 0};
        c->basis.upload(hb, s);
        CUDA_CHECK(cudaStreamSynchronize(s));   // hb goes out of scope
    }
    c->precond.alloc((size_t)pl.S * precond_stride(pl.K));
    c->error_flag.alloc(1);
    CUDA_CHECK(cudaMemsetAsync(c->error_flag.p, 0, sizeof(int), s));
    c->pin.alloc(kPinScalars + (size_t)(pl.P - 1) * nsum);
    if (c->outputs & 1) for (auto& pbuf : c->panels) pbuf.alloc((size_t)pl.P * Np);
    if (c->outputs & 2) c->decisions.alloc((size_t)(pl.P - 1) * Np);
    c->beta_override_set.assign(pl.P - 1, 0);
    c->beta_override.clear();
    c->sims_ready[0] = c->sims_ready[1] = c->backward_done = c->forward_done = false;
    c->fac_set = -1;
    std::memset(&c->tm, 0, sizeof(c->tm));
    CUDA_CHECK(cudaStreamSynchronize(s));
    c->has_problem = true;
}

void simulate(lsmc_ctx* c, int set, const lsmc_rng* rng) {
    require(c->has_problem, LSMC_EINVAL, "lsmc_set_problem has not been called");
    require(set == 0 || set == 1, LSMC_EINVAL, "set must be 0 or 1");
    require(rng != nullptr, LSMC_EINVAL, "rng is null");
    Plan& pl = c->plan;
    const size_t S = pl.S, F = pl.F, N = (size_t)c->N, Np = (size_t)c->Npad;
    cudaStream_t s = c->stream;
    CUDA_CHECK(cudaEventRecord(c->event(set == 0 ? 0 : 4), s));
    // the factor panel is shared by both sets: the valuation set is generated after the backward pass has
    // consumed the regression set, exactly as the reference orders them (LSV.cs:100 vs :346)
    if (set == 1) require(c->backward_done, LSMC_EINVAL, "simulate the valuation set after lsmc_backward");
    if (rng->mode == LSMC_RNG_PANELS) {
        require(rng->spot && rng->factors, LSMC_EINVAL, "LSMC_RNG_PANELS needs spot and factors");
        CUDA_CHECK(cudaMemcpy2DAsync(c->spot[set].p, Np * sizeof(double), rng->spot, N * sizeof(double), N * sizeof(double), S, cudaMemcpyHostToDevice, s));
        CUDA_CHECK(cudaMemcpy2DAsync(c->fac.p, Np * sizeof(double), rng->factors, N * sizeof(double), N * sizeof(double), S * F, cudaMemcpyHostToDevice, s));
    } else {
        SimParams sp{};
        sp.N = c->N; sp.Npad = c->Npad; sp.sim_offset = pl.problem.pb.sim_offset;
        sp.S = pl.S; sp.F = pl.F; sp.mode = rng->mode; sp.antithetic = rng->antithetic; sp.seed = rng->seed;
        if (rng->mode == LSMC_RNG_NORMALS) {
            require(rng->normals != nullptr, LSMC_EINVAL, "LSMC_RNG_NORMALS needs normals");
            c->normals_stage.alloc(S * F * N);
            CUDA_CHECK(cudaMemcpyAsync(c->normals_stage.p, rng->normals, S * F * N * sizeof(double), cudaMemcpyHostToDevice, s));
            sp.normals = c->normals_stage.p;
        } else if (rng->mode == LSMC_RNG_MT19937 || rng->mode == LSMC_RNG_MT19937_CONTINUE) {
            // the reference's seeded stream: one serial generator walks ALL paths of the valuation (every rank of a sharded
            // run generates the whole run and keeps its own columns), then the draws are injected like LSMC_RNG_NORMALS
            if (rng->mode == LSMC_RNG_MT19937) c->mt.seed((uint32_t)rng->seed);
            else require(c->mt.seeded(), LSMC_EINVAL, "LSMC_RNG_MT19937_CONTINUE needs an earlier LSMC_RNG_MT19937 simulation on this context");
            c->mt_normals.resize(S * F * N);
            mt19937_path_normals(c->mt, rng->antithetic != 0, c->Nglobal, pl.problem.pb.sim_offset, c->N, (int)S, (int)F, c->mt_normals.data());
            c->normals_stage.alloc(S * F * N);
            CUDA_CHECK(cudaMemcpyAsync(c->normals_stage.p, c->mt_normals.data(), S * F * N * sizeof(double), cudaMemcpyHostToDevice, s));
            CUDA_CHECK(cudaStreamSynchronize(s));   // pageable source
            sp.mode = LSMC_RNG_NORMALS;
            sp.normals = c->normals_stage.p;
        } else require(rng->mode == LSMC_RNG_PHILOX, LSMC_EINVAL, "unknown rng mode");
        sp.decay = c->s_decay.p; sp.chol = c->s_chol.p; sp.vol = c->s_vol.p; sp.drift = c->s_drift.p; sp.fwd = c->s_fwd.p;
        sp.spot = c->spot[set].p; sp.fac = c->fac.p;
        launch_simulate(sp, s);
        CUDA_CHECK(cudaGetLastError());
        c->tm.kernel_launches++;
    }
    if (set == 0) {
        // regression preconditioners for every simulated period, from the first paths of global rank 0
        PrecondParams pq{};
        pq.N = c->N; pq.Npad = c->Npad; pq.S = pl.S; pq.K = pl.K; pq.F = pl.F; pq.nsub = precond_rows(pl.K, c->N);
        pq.spot = c->spot[0].p; pq.fac = c->fac.p; pq.basis = c->basis.p; pq.out = c->precond.p;
        launch_precond(pq, s);
        CUDA_CHECK(cudaGetLastError());
        c->tm.kernel_launches++;
        c->broadcast0(c->precond.p, (size_t)pl.S * precond_stride(pl.K));
    }
    HOST_TRACE(c, set == 0 ? "simulate 0: launched" : "simulate 1: launched");
    if (set == 1) c->valuation_antithetic = ((rng->mode == LSMC_RNG_PHILOX || rng->mode == LSMC_RNG_MT19937 || rng->mode == LSMC_RNG_MT19937_CONTINUE) && rng->antithetic) ? 1 : 0;
    CUDA_CHECK(cudaEventRecord(c->event(set == 0 ? 1 : 5), s));
    c->sims_ready[set] = true;
    c->fac_set = set;
    if (set == 0) { c->backward_done = false; }
    c->forward_done = false;
}

// Backward induction (LSV.cs:104-341): terminal values, then one fused sweep + reduce + solve per period.
void backward(lsmc_ctx* c, lsmc_progress_cb cb, void* user) {
    require(c->has_problem && c->sims_ready[0], LSMC_EINVAL, "simulate the regression set before lsmc_backward");
    require(c->fac_set == 0, LSMC_EINVAL, "the regression factors were overwritten by the valuation set: simulate the regression set again before lsmc_backward");
    Plan& pl = c->plan;
    const lsmc_problem& pb = pl.problem.pb;
    const int P = pl.P, s0 = pl.s0, F = pl.F, D = pl.max_dec;
    const size_t Np = (size_t)c->Npad;
    cudaStream_t s = c->stream;
    SweepConfig cfg{c->TP, c->MT, c->sweep_blocks, pl.basis_rows.n_scratch, c->nbuf, (int)pl.basis_rows.ops.size()};
    CUDA_CHECK(cudaEventRecord(c->event(2), s));
    c->tm.backward_sweep_launches = 0;
    c->tm.sweep_bytes = 0.0;
    size_t ev_base = 8;
    int cur = 0;
    double progress = 0.0;
    const double back_step = 0.66 / (P - 1);   // BackwardPcntTime (LSV.cs:46,161)

    auto regressors = [&](int period, const double*& spot, const double*& fac) {
        const int j = period - s0;
        spot = (j >= 0) ? c->spot[0].p + (size_t)j * Np : nullptr;
        fac = (j >= 0) ? c->fac.p + (size_t)j * F * Np : nullptr;
    };

    for (int p = P - 1; p >= 0; p--) {
        SweepParams q{};
        q.N = c->N; q.Npad = c->Npad;
        q.G = pl.grid_size(p);
        q.K = pl.K; q.F = F; q.max_dec = D;
        q.ldb = c->ldb;
        q.basis = c->basis.p;
        q.partial = c->partial.p;
        // Programmatic dependent launch, per edge of the loop (measured on one B200, bench.py, us per period outside the sweep):
        //   tail -> sweep (the sweep's set-up overlaps the tail): C3 19.7 -> 9.6, C3 at 125k paths 19.3 -> 8.7 — kept for problems on
        //   the pipelined sweep; C4 (general sweep, 190 KB of shared memory per CTA) 90.8 -> 107.4 — off there;
        //   sweep -> tail (the tail's CTAs start as the sweep's CTAs leave): C3 9.6 -> 17.1, C4 107 -> 166 — off everywhere.
        // LSMC_PDL_EDGES (development): bit 0 tail -> sweep, bit 1 sweep -> tail, bit 2 also with the general sweep.
        static const int pdl_edges = [] { const char* e = getenv("LSMC_PDL_EDGES"); return e ? atoi(e) : 1; }();
        const bool pdl_on = c->use_pdl && pdl_enabled() && (c->TP4 != 0 || (pdl_edges & 4));
        q.pdl = (pdl_on && (pdl_edges & 1)) ? 1 : 0;
        const int pdl_tail = (pdl_on && (pdl_edges & 2)) ? 1 : 0;
        const bool is_current = (p == 0 && s0 == 1);
        double* vout = c->keep_values ? c->Vall.p + (size_t)p * pl.max_grid * Np : c->V(cur ^ 1);
        const double* vnext = c->keep_values ? c->Vall.p + (size_t)(p + 1) * pl.max_grid * Np : c->V(cur);
        q.Vout = vout;
        if (p == P - 1) {
            q.mode = 1; q.Gn = 0;
            q.grid = c->grid.p + pl.grid_offset[p];
            q.must_be_empty = pb.must_be_empty_at_end;
            q.tc0 = pb.terminal_coef[0]; q.tc1 = pb.terminal_coef[1]; q.tc2 = pb.terminal_coef[2]; q.tc3 = pb.terminal_coef[3];
        } else {
            q.mode = 0; q.Gn = pl.grid_size(p + 1);
            q.Vnext = vnext;
            q.cont_mode = is_current ? 1 : 0;
            q.beta = c->beta_all.p + (size_t)p * c->beta_stride();
            q.df = pb.settle_df[p];
            const size_t row = pl.tables.row_offset[p];
            q.t_ndec = c->t_ndec.p + row; q.t_invcost = c->t_invcost.p + row;
            q.t_vol = c->t_vol.p + row * D; q.t_cost = c->t_cost.p + row * D; q.t_cons = c->t_cons.p + row * D;
            q.t_wlo = c->t_wlo.p + row * D; q.t_whi = c->t_whi.p + row * D; q.t_lo = c->t_lo.p + row * D; q.t_hi = c->t_hi.p + row * D;
        }
        q.spot_is_const = is_current ? 1 : 0;
        q.spot_const = pb.forward[p];
        regressors(p, q.spot_p, q.fac_p);
        // normal equations for the regression of period p-1 (ones-only when p-1 is the deterministic current period;
        // at p == 0 the ones column yields the backward NPV, LSV.cs:589)
        if (p == 0) q.accum_mode = 2;
        else {
            q.accum_mode = (p - 1 == 0 && s0 == 1) ? 2 : 1;
            regressors(p - 1, q.spot_m, q.fac_m);
            if (q.accum_mode == 1) q.precond = c->precond.p + (size_t)(p - 1 - s0) * precond_stride(pl.K);
        }
        if (g_trace_host) {
            static thread_local double last = 0.0;
            const double nowt = std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
            if (nowt - last > 0.05) fprintf(stderr, "[rank %d %.6f] backward: period %d reached after %.3f s\n", c->rank, nowt, p, nowt - last);
            last = nowt;
        }
        if (c->time_kernels) CUDA_CHECK(cudaEventRecord(c->event(ev_base), s));
        const bool use4 = c->TP4 != 0 && sweep4_covers(q);
        DevBuf<long long> trace_buf;
        const char* trace_path = getenv("LSMC_TRACE");   // development: phase timestamps of one mid-horizon sweep
        const bool trace_now = trace_path && p == P / 2;
        if (trace_now) { trace_buf.alloc(16 * 64 * 8); CUDA_CHECK(cudaMemsetAsync(trace_buf.p, 0, 16 * 64 * 8 * sizeof(long long), s)); q.trace = trace_buf.p; }
        const int n_parts = use4 ? c->sweep4_blocks : c->sweep_blocks;
        if (use4) CUDA_CHECK(launch_sweep4(q, c->TP4, (int)pl.basis.ops.size(), c->sweep4_blocks, s));
        else CUDA_CHECK(launch_sweep(q, cfg, s));
        if (c->time_kernels) { CUDA_CHECK(cudaEventRecord(c->event(ev_base + 1), s)); ev_base += 2; }
        if (trace_now) {
            std::vector<long long> h(16 * 64 * 8);
            CUDA_CHECK(cudaMemcpyAsync(h.data(), trace_buf.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaStreamSynchronize(s));
            if (FILE* f = fopen(trace_path, "wb")) { fwrite(h.data(), sizeof(long long), h.size(), f); fclose(f); }
        }
        c->tm.backward_sweep_launches++;
        c->tm.kernel_launches++;
        c->tm.sweep_bytes += (double)c->N * 8.0 * ((double)q.G + q.Gn + (q.mode == 0 && !is_current ? F + 1 : (q.mode == 1 ? 1 : 0)) + (q.accum_mode == 1 ? F + 1 : 0));

        const int n_elem = 8 * c->MT * q.ncols;
        SolveParams sp{};
        sp.reduced = c->reduced.p; sp.ncols = q.ncols; sp.K = pl.K; sp.G = q.G; sp.NT = (q.G + 7) / 8; sp.ldb = c->ldb;
        sp.accum_mode = q.accum_mode; sp.n_global = c->Nglobal; sp.pivot_tol = 1e-13;
        sp.gram_lower_blocks = use4 ? 0 : 1;
        sp.precond = q.accum_mode == 1 ? q.precond : nullptr;
        sp.beta = p > 0 ? c->beta_all.p + (size_t)(p - 1) * c->beta_stride() : nullptr;
        if (c->use_tail && c->peer.ready) {
            // one kernel: reduction of the partials, exchange between the ranks over peer memory, solve spread over the CTAs
            TailParams tp{};
            tp.solve = sp; tp.partial = c->partial.p; tp.n_parts = n_parts; tp.n_elem = n_elem;
            tp.do_solve = p > 0 ? 1 : 0;
            tp.total_out = c->reduced.p; tp.n_out = 1;   // element 0 at p == 0: sum of the start-inventory values (backward NPV)
            tp.peers = c->peer.view; tp.error_flag = c->error_flag.p; tp.pdl = pdl_tail;
            const int blocks = tail_blocks(n_elem, q.G, c->num_sms);
            c->peer.next(blocks, &tp.parity, &tp.target, &tp.local_target);
            static const bool tail_trace = getenv("LSMC_TAIL_TRACE") != nullptr;   // development: phase stamps of one mid-horizon tail
            DevBuf<long long> tail_tr;
            if (tail_trace && p == P / 2) { tail_tr.alloc(8); CUDA_CHECK(cudaMemsetAsync(tail_tr.p, 0, 64, s)); tp.trace = tail_tr.p; }
            launch_tail(tp, blocks, s);
            c->tm.kernel_launches++;
            if (tp.trace) {
                long long h[8];
                CUDA_CHECK(cudaMemcpyAsync(h, tail_tr.p, 64, cudaMemcpyDeviceToHost, s));
                CUDA_CHECK(cudaStreamSynchronize(s));
                fprintf(stderr, "tail trace (ns): reduce %lld, barrier %lld, stage %lld, factor %lld, solve+store %lld, total %lld\n", h[1] - h[0], h[2] - h[1],
                        h[3] - h[2], h[4] - h[3], h[6] - h[4], h[6] - h[0]);
            }
        } else {
            launch_reduce_partials(c->partial.p, n_parts, n_elem, c->reduced.p, s);
            c->tm.kernel_launches++;
            c->allreduce(c->reduced.p, (size_t)n_elem);
            if (p > 0) { launch_solve(sp, s); c->tm.kernel_launches++; }
        }
        if (p > 0) {
            if (c->beta_override_set[p - 1])
                CUDA_CHECK(cudaMemcpyAsync(sp.beta, c->beta_override.data() + (size_t)(p - 1) * c->beta_stride(),
                                           c->beta_stride() * sizeof(double), cudaMemcpyHostToDevice, s));
        } else {
            CUDA_CHECK(cudaMemcpyAsync(c->pin.p, c->reduced.p, sizeof(double), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaStreamSynchronize(s));
            c->backward_npv = c->pin.p[0] / (double)c->Nglobal;
        }
        CUDA_CHECK(cudaGetLastError());
        cur ^= 1;
        if (p < P - 1) {   // one progress tick per decision period (LSV.cs:337-339)
            progress += back_step;
            if (cb) {
                CUDA_CHECK(cudaStreamSynchronize(s));
                if (cb(user, progress) != 0) {
                    if (c->group) c->group->cancelled.store(1);
                    throw Error(LSMC_ECANCELLED, "The operation was canceled.");
                }
            }
            if (c->group && c->group->cancelled.load()) throw Error(LSMC_ECANCELLED, "The operation was canceled.");
        }
    }
    HOST_TRACE(c, "backward: all periods launched");
    CUDA_CHECK(cudaEventRecord(c->event(3), s));
    check_device_error_flag(c);   // synchronises the stream; a timed-out exchange surfaces here
    HOST_TRACE(c, "backward: done");
    float ms = 0.f;
    c->tm.backward_sweep_kernel_ms = 0.0;
    for (size_t e = 8; e < ev_base; e += 2) { CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev[e], c->ev[e + 1])); c->tm.backward_sweep_kernel_ms += ms; }
    CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev[2], c->ev[3])); c->tm.backward_ms = ms;
    CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1])); c->tm.regression_sim_ms = ms;
    c->backward_done = true;
}

// Forward simulation (LSV.cs:344-579).
void forward(lsmc_ctx* c, lsmc_progress_cb cb, void* user) {
    require(c->backward_done && c->sims_ready[1] && c->fac_set == 1, LSMC_EINVAL, "run lsmc_backward and simulate the valuation set before lsmc_forward");
    Plan& pl = c->plan;
    const lsmc_problem& pb = pl.problem.pb;
    const int P = pl.P, s0 = pl.s0, F = pl.F, K = pl.K, nsum = kFwdSums + K;
    const size_t Np = (size_t)c->Npad;
    cudaStream_t s = c->stream;
    HOST_TRACE(c, "forward: begin");
    CUDA_CHECK(cudaEventRecord(c->event(6), s));
    launch_fill(c->inventory.p, c->Npad, pb.start_inventory, s);   // LSV.cs:362-364
    CUDA_CHECK(cudaMemsetAsync(c->pv_by_sim.p, 0, Np * sizeof(double), s));
    CUDA_CHECK(cudaMemsetAsync(c->error_flag.p, 0, sizeof(int), s));
    const bool panels = (c->outputs & 1) != 0;
    double progress = 0.66;
    const double fwd_step = (1.0 - 0.66) / P;   // LSV.cs:371
    std::vector<ForwardParams> hq((size_t)(P - 1));
    for (int p = 0; p < P - 1; p++) {
        ForwardParams& q = hq[p];
        q = ForwardParams{};
        q.N = c->N; q.Npad = c->Npad; q.p = p; q.Gn = pl.grid_size(p + 1); q.K = K; q.F = F; q.max_dec = pl.max_dec; q.extra = pb.extra_decisions;
        const bool is_current = (p == 0 && s0 == 1);
        q.cont_mode = is_current ? 1 : 0; q.spot_is_const = is_current ? 1 : 0; q.spot_const = pb.forward[p];
        q.df = pb.settle_df[p]; q.tol = pb.numerical_tolerance;
        q.loss_pct = pb.inventory_loss[p]; q.inv_cost = pb.inventory_cost[p]; q.inv_cost_df = pb.inventory_cost_df[p];
        q.inj_cost = pb.inject_cost[p]; q.inj_df = pb.inject_cost_df[p]; q.wd_cost = pb.withdraw_cost[p]; q.wd_df = pb.withdraw_cost_df[p];
        q.cons_inj = pb.consumed_inject[p]; q.cons_wd = pb.consumed_withdraw[p];
        q.min_inv = pb.min_inventory[p]; q.max_inv = pb.max_inventory[p];
        q.slack_lo = range_slack(pb.numerical_tolerance, q.min_inv); q.slack_hi = range_slack(pb.numerical_tolerance, q.max_inv);
        q.next_min = pl.space_min[p + 1]; q.next_max = pl.space_max[p + 1];
        const int cid = pb.constraint_id[p], first = pb.constraint_first[cid];
        q.c_kind = pb.constraint_kind[cid]; q.c_n = pb.constraint_count[cid];
        q.c_x = c->c_x.p + first; q.c_lo = c->c_lo.p + first; q.c_hi = c->c_hi.p + first;
        if (!is_current) { q.spot = c->spot[1].p + (size_t)(p - s0) * Np; q.fac = c->fac.p + (size_t)(p - s0) * F * Np; }
        q.beta = c->beta_all.p + (size_t)p * c->beta_stride(); q.ldb = c->ldb;
        q.next_grid = c->grid.p + pl.grid_offset[p + 1];
        q.basis = c->basis.p;
        q.inventory = c->inventory.p; q.pv_by_sim = c->pv_by_sim.p;
        q.sums = c->fwd_sums.p + (size_t)p * c->fwd_parts * nsum;
        q.error_flag = c->error_flag.p;
        if (panels) {
            q.o_inventory = c->panels[0].p + (size_t)p * Np; q.o_volume = c->panels[1].p + (size_t)p * Np; q.o_consumed = c->panels[2].p + (size_t)p * Np;
            q.o_loss = c->panels[3].p + (size_t)p * Np; q.o_net = c->panels[4].p + (size_t)p * Np; q.o_pv = c->panels[5].p + (size_t)p * Np;
        }
        if (c->outputs & 2) q.o_decision = c->decisions.p + (size_t)p * Np;
    }
    HOST_TRACE(c, "forward: upload params");
    c->fwd_params.upload(hq, s);
    HOST_TRACE(c, "forward: launch");
    Forward2Args fa{};
    fa.params = c->fwd_params.p; fa.K = K; fa.F = F; fa.n_nodes = pl.basis.n_nodes; fa.max_grid = pl.max_grid; fa.ldb = c->ldb;
    fa.grid_pitch = (pl.max_grid + 1) & ~1; fa.n_parts = c->fwd_parts; fa.basis = c->basis.p;
    fa.beta_all = c->beta_all.p; fa.beta_stride = c->beta_stride(); fa.grid_pad = c->grid_pad.p; fa.pillars = c->fwd_pillars.p;
    fa.warp_sums = c->fwd_sums.p;
    auto launch_periods = [&](int p0, int p1) {
        if (c->fwd2) { fa.p0 = p0; fa.p1 = p1; CUDA_CHECK(launch_forward2(fa, K, c->fwd_blocks, s)); }
        else launch_forward(c->fwd_params.p, p0, p1, K, pl.basis.n_nodes, pl.max_grid, c->fwd_blocks, c->fwd_fast, s);
        c->tm.kernel_launches++;
    };
    if (!cb) {   // the whole pass in one launch
        launch_periods(0, P - 1);
    } else {     // one launch per period so that progress / cancellation keep the reference's cadence (LSV.cs:488-490)
        for (int p = 0; p < P - 1; p++) {
            launch_periods(p, p + 1);
            progress += fwd_step;
            CUDA_CHECK(cudaStreamSynchronize(s));
            check_device_error_flag(c);
            if (cb(user, progress) != 0) {
                if (c->group) c->group->cancelled.store(1);
                throw Error(LSMC_ECANCELLED, "The operation was canceled.");
            }
        }
    }
    CUDA_CHECK(cudaGetLastError());
    HOST_TRACE(c, "forward: reduce + allreduce");
    launch_reduce_rows(c->fwd_sums.p, P - 1, c->fwd_parts, nsum, c->fwd_red.p, s);
    c->tm.kernel_launches++;
    c->allreduce(c->fwd_red.p, (size_t)(P - 1) * nsum);
    HOST_TRACE(c, "forward: final");
    // final period + NPV
    FinalParams fq{};
    fq.N = c->N; fq.pass = 0; fq.must_be_empty = pb.must_be_empty_at_end;
    fq.tc0 = pb.terminal_coef[0]; fq.tc1 = pb.terminal_coef[1]; fq.tc2 = pb.terminal_coef[2]; fq.tc3 = pb.terminal_coef[3];
    fq.spot_end = c->spot[0].p + (size_t)(pl.S - 1) * Np;   // regression sims' end prices (LSV.cs:567)
    fq.inventory = c->inventory.p; fq.pv_by_sim = c->pv_by_sim.p; fq.sums = c->fin_sums.p;
    if (panels) {
        fq.o_inventory = c->panels[0].p + (size_t)(P - 1) * Np; fq.o_pv = c->panels[5].p + (size_t)(P - 1) * Np;
        for (int k = 1; k <= 4; k++) CUDA_CHECK(cudaMemsetAsync(c->panels[k].p + (size_t)(P - 1) * Np, 0, Np * sizeof(double), s));
    }
    launch_final(fq, c->fwd_blocks, s);
    launch_reduce_rows(c->fin_sums.p, 1, c->fwd_blocks, 3, c->reduced.p, s);
    c->tm.kernel_launches += 2;
    c->allreduce(c->reduced.p, 3);
    double* fin = c->pin.p + 1;
    CUDA_CHECK(cudaMemcpyAsync(fin, c->reduced.p, 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
    c->h_fwd.resize((size_t)(P - 1) * nsum);
    CUDA_CHECK(cudaMemcpyAsync(c->pin.p + kPinScalars, c->fwd_red.p, c->h_fwd.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    std::memcpy(c->h_fwd.data(), c->pin.p + kPinScalars, c->h_fwd.size() * sizeof(double));
    check_device_error_flag(c);
    c->final_inventory_sum = fin[0]; c->terminal_pv_sum = fin[1];
    c->npv = fin[2] / (double)c->Nglobal;   // LSV.cs:584
    // standard error of the NPV (pair means when the valuation draws are antithetic)
    // pair means only when EVERY path of the valuation has its partner: decided on the global count, so that all ranks of a
    // sharded valuation agree (shards start on even paths and are even except the last: checked in set_problem)
    const bool paired = c->valuation_antithetic && (c->Nglobal % 2 == 0) && (c->N % 2 == 0) && (pb.sim_offset % 2 == 0);
    fq.pass = 1; fq.paired = paired ? 1 : 0; fq.mean = c->npv;
    launch_final(fq, c->fwd_blocks, s);
    launch_reduce_rows(c->fin_sums.p, 1, c->fwd_blocks, 3, c->reduced.p, s);
    c->tm.kernel_launches += 2;
    c->allreduce(c->reduced.p, 3);
    double* ss = c->pin.p + 4;
    CUDA_CHECK(cudaMemcpyAsync(ss, c->reduced.p, 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaEventRecord(c->event(7), s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    const double m = (double)(paired ? c->Nglobal / 2 : c->Nglobal);
    c->npv_se = m > 1 ? std::sqrt(ss[0] / (m - 1) / m) : 0.0;
    float ms = 0.f;
    CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev[4], c->ev[5])); c->tm.valuation_sim_ms = ms;
    CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev[6], c->ev[7])); c->tm.forward_ms = ms;
    c->tm.total_ms = c->tm.regression_sim_ms + c->tm.backward_ms + c->tm.valuation_sim_ms + c->tm.forward_ms;
    c->tm.span_ms = 0.0;
    if (c->ev.size() > 7 && c->ev[0] && c->ev[7] && cudaEventElapsedTime(&ms, c->ev[0], c->ev[7]) == cudaSuccess && ms > 0.f) c->tm.span_ms = ms;
    else (void)cudaGetLastError();   // regression set injected / simulated before an earlier forward pass: no span
    c->forward_done = true;
    if (cb) cb(user, 1.0);   // LSV.cs:604
}

// Expected profile, deltas (LSV.cs:482-487,593-597) and trigger prices (LSV.cs:492-561) from the per-period sums.
void assemble_summary(lsmc_ctx* c, lsmc_results* r) {
    Plan& pl = c->plan;
    const lsmc_problem& pb = pl.problem.pb;
    const int P = pl.P, K = pl.K, nsum = kFwdSums + K, s0 = pl.s0;
    const double Ng = (double)c->Nglobal;
    const double NaN = std::numeric_limits<double>::quiet_NaN();
    r->npv = c->npv; r->npv_std_error = c->npv_se; r->backward_npv = c->backward_npv;
    if (r->deltas) {
        for (int p = 0; p < P - 1; p++) {
            const double* sm = &c->h_fwd[(size_t)p * nsum];
            const double disc = pb.discount_deltas ? pb.settle_df[p] : 1.0;
            r->deltas[p] = (sm[6] / pb.forward[p] / Ng) * disc;
        }
        r->deltas[P - 1] = 0.0;
    }
    if (r->profile) {
        for (int p = 0; p < P - 1; p++)
            for (int k = 0; k < 6; k++) r->profile[(size_t)p * 6 + k] = c->h_fwd[(size_t)p * nsum + k] / Ng;
        double* last = r->profile + (size_t)(P - 1) * 6;
        last[0] = c->final_inventory_sum / Ng; last[1] = last[2] = last[3] = last[4] = 0.0;
        last[5] = pb.must_be_empty_at_end ? 0.0 : c->terminal_pv_sum / Ng;
    }
    const bool want_trig = r->trigger_summary || r->trigger_inject || r->trigger_withdraw;
    if (!want_trig) return;
    const int nv = LSMC_NUM_TRIGGER_VOLUMES;
    if (r->trigger_summary) std::fill(r->trigger_summary, r->trigger_summary + (size_t)(P - 1) * 4, NaN);
    if (r->trigger_inject) std::fill(r->trigger_inject, r->trigger_inject + (size_t)(P - 1) * nv * 2, NaN);
    if (r->trigger_withdraw) std::fill(r->trigger_withdraw, r->trigger_withdraw + (size_t)(P - 1) * nv * 2, NaN);
    std::vector<double> beta(c->beta_all.n);
    CUDA_CHECK(cudaMemcpy(beta.data(), c->beta_all.p, beta.size() * sizeof(double), cudaMemcpyDeviceToHost));
    std::vector<double> ds(pl.max_dec), xbar(K);
    for (int p = 0; p < P - 1; p++) {
        const double* sm = &c->h_fwd[(size_t)p * nsum];
        const bool is_current = (p == 0 && s0 == 1);
        const double inv = sm[0] / Ng;   // expected inventory
        for (int k = 0; k < K; k++) xbar[k] = sm[kFwdSums + k] / Ng;
        const double* bp = beta.data() + (size_t)p * c->beta_stride();
        const double* next = pl.grid_of(p + 1);
        const int Gn = pl.grid_size(p + 1);
        // mean over sims of the fitted continuation value at grid row l = xbar . beta_l
        auto mean_fit = [&](int l) {
            if (is_current) return bp[(size_t)l * c->ldb];
            double acc = 0.0;
            for (int k = 0; k < K; k++) acc += xbar[k] * bp[(size_t)l * c->ldb + k];
            return acc;
        };
        auto avg_cont = [&](double after) {   // AverageContinuationValue (LSV.cs:700-720)
            int lo, hi;
            bisect_grid(next, Gn, after, pb.numerical_tolerance, lo, hi);
            if (lo == hi) return mean_fit(lo);
            const double wl = (next[hi] - after) / (next[hi] - next[lo]);
            return mean_fit(lo) * wl + mean_fit(hi) * (1.0 - wl);
        };
        const double loss = pb.inventory_loss[p] * inv;
        double rmin, rmax;
        inject_withdraw_range(pb, p, inv, rmin, rmax);
        const int nd = bang_bang_decisions(rmin, rmax, inv, loss, pl.space_min[p + 1], pl.space_max[p + 1], pb.numerical_tolerance, pb.extra_decisions, ds.data());
        const double df = pb.settle_df[p];
        auto price_of = [&](double vol, double alt_cont, double alt_vol, double alt_cost, double alt_cons) {   // CalcTriggerPrice (LSV.cs:619-638)
            const double cont_change = avg_cont(inv + vol - loss) - alt_cont;
            const double cost_change = decision_cost_npv(pb, p, vol) - alt_cost;
            const double cons_change = decision_consumed(pb, p, vol) - alt_cons;
            return (cont_change - cost_change) / (df * ((vol - alt_vol) + cons_change));
        };
        double max_inj = ds[0], max_wd = ds[0];
        for (int d = 1; d < nd; d++) { max_inj = std::max(max_inj, ds[d]); max_wd = std::min(max_wd, ds[d]); }
        if (max_inj > 0) {
            double alt = std::numeric_limits<double>::infinity();
            for (int d = 0; d < nd; d++) if (ds[d] >= 0 && ds[d] < alt) alt = ds[d];
            if (max_inj > alt) {
                const double alt_cont = avg_cont(inv + alt - loss), alt_cost = decision_cost_npv(pb, p, alt), alt_cons = decision_consumed(pb, p, alt);
                const double inc = (max_inj - alt) / nv;
                double last_price = NaN;
                for (int i = 0; i < nv; i++) {
                    const double vol = (i == nv - 1) ? max_inj : alt + (i + 1) * inc;   // LSV.cs:640-649
                    last_price = price_of(vol, alt_cont, alt, alt_cost, alt_cons);
                    if (r->trigger_inject) { r->trigger_inject[((size_t)p * nv + i) * 2] = vol; r->trigger_inject[((size_t)p * nv + i) * 2 + 1] = last_price; }
                }
                if (r->trigger_summary) { r->trigger_summary[(size_t)p * 4] = max_inj; r->trigger_summary[(size_t)p * 4 + 1] = last_price; }
            }
        }
        if (max_wd < 0) {
            double alt = -std::numeric_limits<double>::infinity();
            for (int d = 0; d < nd; d++) if (ds[d] <= 0 && ds[d] > alt) alt = ds[d];
            if (max_wd < alt) {
                const double alt_cont = avg_cont(inv + alt - loss), alt_cost = decision_cost_npv(pb, p, alt), alt_cons = decision_consumed(pb, p, alt);
                const double inc = (alt - max_wd) / nv;
                for (int out = 0; out < nv; out++) {   // volumes max_wd + i*inc, listed in reverse (LSV.cs:545,651-659)
                    const int i = nv - 1 - out;
                    const double vol = max_wd + i * inc;
                    const double price = price_of(vol, alt_cont, alt, alt_cost, alt_cons);
                    if (r->trigger_withdraw) { r->trigger_withdraw[((size_t)p * nv + out) * 2] = vol; r->trigger_withdraw[((size_t)p * nv + out) * 2 + 1] = price; }
                    if (out == 0 && r->trigger_summary) { r->trigger_summary[(size_t)p * 4 + 2] = max_wd; r->trigger_summary[(size_t)p * 4 + 3] = price; }
                }
            }
        }
    }
}

// Per-sim outputs: rows of N doubles into the caller's [rows][pitch] buffer at column out_col0 (a rank of a multi-GPU
// valuation writes its own block of columns; a single context has pitch = N, column 0).
//
// Export path (the eight [P][N] panels are 23 GB at C3).  A destination in page-locked memory
// (lsmc_host_alloc, cudaHostAlloc, cudaHostRegister) is written by the copy engine directly: every panel is enqueued on the
// context's stream and the call synchronises once.  A pageable destination goes through two page-locked staging chunks:
// the copy engine fills one while the calling thread empties the other into the caller's array.
constexpr size_t kExportChunkBytes = (size_t)32 << 20;

bool is_page_locked(const void* p) {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return attr.type == cudaMemoryTypeHost;
}

void copy_panel(lsmc_ctx* c, double* dst, const double* src, size_t rows) {
    if (!dst || rows == 0) return;
    const size_t pitch = c->out_pitch > 0 ? (size_t)c->out_pitch : (size_t)c->N, n = (size_t)c->N;
    const size_t src_pitch = (size_t)c->Npad * sizeof(double), width = n * sizeof(double);
    cudaStream_t s = c->stream;
    dst += c->out_col0;
    if (is_page_locked(dst)) {
        CUDA_CHECK(cudaMemcpy2DAsync(dst, pitch * sizeof(double), src, src_pitch, width, rows, cudaMemcpyDeviceToHost, s));
        c->export_pending = true;
        return;
    }
    if (rows * width <= ((size_t)1 << 20)) {   // small: not worth the staging
        CUDA_CHECK(cudaMemcpy2DAsync(dst, pitch * sizeof(double), src, src_pitch, width, rows, cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        return;
    }
    c->export_stage.alloc(2 * kExportChunkBytes / sizeof(double));
    const size_t rows_per_chunk = std::max<size_t>(1, kExportChunkBytes / width);
    if (rows_per_chunk * width > kExportChunkBytes) {   // one row is larger than a chunk (> 4M paths): plain copy
        CUDA_CHECK(cudaMemcpy2DAsync(dst, pitch * sizeof(double), src, src_pitch, width, rows, cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        return;
    }
    for (auto& e : c->export_ev) if (!e) CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    cudaEvent_t* done = c->export_ev;
    double* stage[2] = {c->export_stage.p, c->export_stage.p + kExportChunkBytes / sizeof(double)};
    const size_t n_chunks = (rows + rows_per_chunk - 1) / rows_per_chunk;
    auto issue = [&](size_t k) {
        const size_t r0 = k * rows_per_chunk, nr = std::min(rows_per_chunk, rows - r0);
        CUDA_CHECK(cudaMemcpy2DAsync(stage[k & 1], width, src + r0 * (size_t)c->Npad, src_pitch, width, nr, cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaEventRecord(done[k & 1], s));
    };
    issue(0);
    for (size_t k = 0; k < n_chunks; k++) {
        if (k + 1 < n_chunks) issue(k + 1);
        CUDA_CHECK(cudaEventSynchronize(done[k & 1]));
        const size_t r0 = k * rows_per_chunk, nr = std::min(rows_per_chunk, rows - r0);
        for (size_t r = 0; r < nr; r++) std::memcpy(dst + (r0 + r) * pitch, stage[k & 1] + r * n, width);
    }
}

void get_results(lsmc_ctx* c, lsmc_results* r) {
    require(r != nullptr, LSMC_EINVAL, "results is null");
    require(r->struct_size == (int32_t)sizeof(lsmc_results), LSMC_EINVAL, "lsmc_results.struct_size mismatch");
    require(c->forward_done, LSMC_EINVAL, "lsmc_forward has not completed");
    Plan& pl = c->plan;
    assemble_summary(c, r);
    copy_panel(c, r->pv_by_sim, c->pv_by_sim.p, 1);
    copy_panel(c, r->spot_regress, c->spot[0].p, pl.S);
    copy_panel(c, r->spot_valuation, c->spot[1].p, pl.S);
    double* outs[6] = {r->inventory, r->inject_withdraw, r->consumed, r->loss, r->net_volume, r->pv};
    for (int k = 0; k < 6; k++) {
        if (!outs[k]) continue;
        require((c->outputs & 1) != 0, LSMC_EINVAL, "per-sim panels were not requested before lsmc_forward (lsmc_request_outputs)");
        copy_panel(c, outs[k], c->panels[k].p, pl.P);
    }
    if (r->decision_index) {
        require((c->outputs & 2) != 0, LSMC_EINVAL, "decision_index was not requested before lsmc_forward (lsmc_request_outputs)");
        const size_t pitch = c->out_pitch > 0 ? (size_t)c->out_pitch : (size_t)c->N;
        CUDA_CHECK(cudaMemcpy2D(r->decision_index + c->out_col0, pitch, c->decisions.p, (size_t)c->Npad, (size_t)c->N, pl.P - 1, cudaMemcpyDeviceToHost));
    }
    if (c->export_pending) { CUDA_CHECK(cudaStreamSynchronize(c->stream)); c->export_pending = false; }
}

}  // namespace

// ================================================================================================================
extern "C" {

const char* lsmc_version(void) { return "lsmc_b200 0.1.0 (sm_100a)"; }

int lsmc_create(int device, lsmc_ctx** out) {
    if (!out) { g_create_error = "out is null"; return LSMC_EINVAL; }
    *out = nullptr;
    lsmc_ctx* c = nullptr;
    int rc = guarded(nullptr, [&] {
        int count = 0;
        CUDA_CHECK(cudaGetDeviceCount(&count));
        require(device >= 0 && device < count, LSMC_ECUDA, "no such CUDA device");
        CUDA_CHECK(cudaSetDevice(device));
        cudaDeviceProp prop;
        CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
        require(prop.major >= 10, LSMC_ECUDA, "this library is built for sm_100a (B200) only");
        c = new lsmc_ctx();
        c->device = device;
        c->num_sms = prop.multiProcessorCount;
        CUDA_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    });
    if (rc != LSMC_OK) { delete c; return rc; }
    *out = c;
    return LSMC_OK;
}

void lsmc_destroy(lsmc_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    delete ctx;
}

const char* lsmc_last_error(const lsmc_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int lsmc_comm_unique_id(void* out128) {
    return guarded(nullptr, [&] {
        std::string err;
        require(out128 != nullptr, LSMC_EINVAL, "out128 is null");
        if (!g_nccl.load(err)) throw Error(LSMC_ECUDA, err);
        int rc = g_nccl.GetUniqueId(out128);
        if (rc != 0) throw Error(LSMC_ECUDA, "ncclGetUniqueId failed");
    });
}

int lsmc_comm_init(lsmc_ctx* ctx, int rank, int world, const void* unique_id128) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] {
        require(world >= 1 && rank >= 0 && rank < world, LSMC_EINVAL, "bad rank/world");
        ctx->rank = rank; ctx->world = world;
        if (world == 1) return;
        require(unique_id128 != nullptr, LSMC_EINVAL, "unique id is null");
        std::string err;
        if (!g_nccl.load(err)) throw Error(LSMC_ECUDA, err);
        NcclId id;
        std::memcpy(id.bytes, unique_id128, 128);
        typedef int (*init_fn)(void**, int, NcclId, int);
        int rc = ((init_fn)g_nccl.CommInitRank)(&ctx->comm, world, id, rank);
        if (rc != 0) throw Error(LSMC_ECUDA, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"));
        map_peers_over_ipc(ctx);
    });
}

static int calculate_impl(lsmc_ctx* ctx, const lsmc_problem* problem, const lsmc_rng* regression_rng, const lsmc_rng* valuation_rng,
                          lsmc_progress_cb progress, void* user, lsmc_results* results, const std::function<bool()>& after_setup);

int lsmc_comm_init_local(lsmc_ctx** ctxs, int n) {
    if (!ctxs || n < 1 || n > kMaxPeers) return LSMC_EINVAL;
    for (int r = 0; r < n; r++) if (!ctxs[r]) return LSMC_EINVAL;
    return guarded(ctxs[0], [&] {
        auto group = std::make_shared<LocalGroup>();
        for (int r = 0; r < n; r++) {
            lsmc_ctx* c = ctxs[r];
            require(c->comm == nullptr, LSMC_EINVAL, "context already belongs to an NCCL communicator");
            CUDA_CHECK(cudaSetDevice(c->device));
            c->rank = r; c->world = n; c->group = n > 1 ? group : nullptr;
            for (int o = 0; o < n; o++) if (o != r && ctxs[o]->device == c->device) c->use_pdl = false;
            if (n == 1) { c->peer.self_only(); continue; }
            c->peer.release();
            c->peer.allocate(n);
            for (int o = 0; o < n; o++) {   // direct loads/stores between distinct devices
                if (ctxs[o]->device == c->device) continue;
                int can = 0;
                CUDA_CHECK(cudaDeviceCanAccessPeer(&can, c->device, ctxs[o]->device));
                require(can != 0, LSMC_ECUDA, "the GPUs of the group cannot access each other's memory");
                cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[o]->device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CUDA_CHECK(e);
                (void)cudaGetLastError();
            }
        }
        if (n == 1) return;
        for (int r = 0; r < n; r++) {
            lsmc_ctx* c = ctxs[r];
            PeerView v{};
            v.rank = r; v.world = n;
            for (int o = 0; o < n; o++) { v.slots[o] = ctxs[o]->peer.buf; v.counter[o] = ctxs[o]->peer.counter_of(ctxs[o]->peer.buf, n); }
            v.ticket = ctxs[r]->peer.ticket_of(ctxs[r]->peer.buf, n);
            c->peer.view = v; c->peer.ready = true;
        }
        CUDA_CHECK(cudaSetDevice(ctxs[0]->device));
    });
}

// Even split of the global paths over the ranks on pair boundaries (antithetic partners 2m, 2m+1 stay on one rank):
// the same rule as storage_b200/sharding.py shard_paths.
static void shard_of(int64_t n_global, int world, int rank, int64_t* offset, int64_t* count) {
    const int64_t pairs = (n_global + 1) / 2, base = pairs / world, rem = pairs % world;
    const int64_t first = rank * base + std::min<int64_t>(rank, rem), n_pairs = base + (rank < rem ? 1 : 0);
    *offset = std::min<int64_t>(2 * first, n_global);
    *count = std::min<int64_t>(2 * (first + n_pairs), n_global) - *offset;
}

int lsmc_calculate_multi(lsmc_ctx** ctxs, int n, const lsmc_problem* problem, const lsmc_rng* regression_rng,
                         const lsmc_rng* valuation_rng, lsmc_progress_cb progress, void* user, lsmc_results* results) {
    if (!ctxs || n < 1 || n > kMaxPeers || !ctxs[0]) return LSMC_EINVAL;
    if (n == 1) return lsmc_calculate(ctxs[0], problem, regression_rng, valuation_rng, progress, user, results);
    lsmc_ctx* c0 = ctxs[0];
    if (!problem || !regression_rng || !valuation_rng || !results) { c0->err = "null argument"; return LSMC_EINVAL; }
    for (int r = 0; r < n; r++)
        if (!ctxs[r] || ctxs[r]->world != n || ctxs[r]->rank != r || !ctxs[r]->peer.ready) { c0->err = "call lsmc_comm_init_local on these contexts first"; return LSMC_EINVAL; }
    const int64_t Ng = problem->n_sims;
    if (problem->n_sims_global != 0 && problem->n_sims_global != Ng) { c0->err = "lsmc_calculate_multi takes the whole problem: n_sims_global must be 0"; return LSMC_EINVAL; }
    if (Ng < 2 * n) { c0->err = "fewer than one pair of paths per GPU"; return LSMC_EINVAL; }
    const int S = problem->n_periods - (problem->first_period_is_current ? 1 : 0), F = problem->n_factors;
    if (c0->group) c0->group->cancelled.store(0);
    std::vector<int> status(n, LSMC_OK);
    auto run_rank = [&](int r) {
        lsmc_ctx* c = ctxs[r];
        int64_t off = 0, cnt = 0;
        shard_of(Ng, n, r, &off, &cnt);
        lsmc_problem pb = *problem;
        pb.n_sims = cnt; pb.n_sims_global = Ng; pb.sim_offset = problem->sim_offset + off;
        // injected draws / panels: this rank's columns, repacked contiguously
        std::vector<double> stage[2][3];
        lsmc_rng rng[2] = {*regression_rng, *valuation_rng};
        for (int k = 0; k < 2; k++) {
            auto slice = [&](const double* src, size_t rows, std::vector<double>& dst) -> const double* {
                if (!src) return nullptr;
                dst.resize(rows * (size_t)cnt);
                for (size_t row = 0; row < rows; row++) std::memcpy(dst.data() + row * cnt, src + row * Ng + off, sizeof(double) * cnt);
                return dst.data();
            };
            if (rng[k].mode == LSMC_RNG_NORMALS) rng[k].normals = slice(rng[k].normals, (size_t)S * F, stage[k][0]);
            if (rng[k].mode == LSMC_RNG_PANELS) { rng[k].spot = slice(rng[k].spot, (size_t)S, stage[k][1]); rng[k].factors = slice(rng[k].factors, (size_t)S * F, stage[k][2]); }
        }
        lsmc_results res = *results;
        if (r != 0) {   // the summary is identical on every rank: rank 0 writes it
            res.deltas = res.profile = res.trigger_summary = res.trigger_inject = res.trigger_withdraw = nullptr;
        }
        c->out_pitch = Ng; c->out_col0 = off;
        // all ranks finish their set-up (allocations) before any of them starts a kernel that waits for the others
        auto after_setup = [&]() -> bool { if (c->group) { c->group->barrier(n); return c->group->cancelled.load() == 0; } return true; };
        status[r] = calculate_impl(c, &pb, &rng[0], &rng[1], r == 0 ? progress : nullptr, user, &res, after_setup);
        c->out_pitch = 0; c->out_col0 = 0;
        if (r == 0) { results->npv = res.npv; results->npv_std_error = res.npv_std_error; results->backward_npv = res.backward_npv; }
        else if (status[r] != LSMC_OK && c->group) c->group->cancelled.store(1);
    };
    std::vector<std::thread> workers;
    for (int r = 1; r < n; r++) workers.emplace_back(run_rank, r);
    run_rank(0);   // the calling thread is rank 0: progress callbacks arrive on it, as in the reference
    for (auto& w : workers) w.join();
    // report the failure that started it, not the cancellations it caused on the other ranks
    int first = -1;
    for (int r = 0; r < n; r++) if (status[r] != LSMC_OK && (first < 0 || (status[first] == LSMC_ECANCELLED && status[r] != LSMC_ECANCELLED))) first = r;
    if (first < 0) return LSMC_OK;
    if (first != 0) c0->err = "GPU " + std::to_string(first) + " of the group: " + ctxs[first]->err;
    return status[first];
}

int lsmc_request_outputs(lsmc_ctx* ctx, uint32_t mask) {
    if (!ctx) return LSMC_EINVAL;
    ctx->outputs = mask;
    return LSMC_OK;
}

int lsmc_set_problem(lsmc_ctx* ctx, const lsmc_problem* problem) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] { set_problem(ctx, problem); });
}

int lsmc_get_layout(const lsmc_ctx* ctx, lsmc_layout* out) {
    if (!ctx || !out || !ctx->has_problem) return LSMC_EINVAL;
    out->n_periods = ctx->plan.P; out->n_sim_periods = ctx->plan.S; out->max_grid = ctx->plan.max_grid;
    out->max_decisions = ctx->plan.max_dec; out->n_sims = ctx->N;
    return LSMC_OK;
}

int lsmc_simulate(lsmc_ctx* ctx, int set, const lsmc_rng* rng) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] { simulate(ctx, set, rng); });
}

int lsmc_backward(lsmc_ctx* ctx, lsmc_progress_cb progress, void* user) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] { backward(ctx, progress, user); });
}

int lsmc_forward(lsmc_ctx* ctx, lsmc_progress_cb progress, void* user) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] { forward(ctx, progress, user); });
}

int lsmc_get_results(lsmc_ctx* ctx, lsmc_results* results) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] { get_results(ctx, results); });
}

// lsmc_calculate, with a hook between the set-up (host plan, every device allocation of the valuation) and the first kernel.
static int calculate_impl(lsmc_ctx* ctx, const lsmc_problem* problem, const lsmc_rng* regression_rng, const lsmc_rng* valuation_rng,
                          lsmc_progress_cb progress, void* user, lsmc_results* results, const std::function<bool()>& after_setup) {
    if (!ctx) return LSMC_EINVAL;
    bool setup_done = false;
    int rc = guarded(ctx, [&] {
        require(results != nullptr, LSMC_EINVAL, "results is null");
        require(regression_rng != nullptr && valuation_rng != nullptr, LSMC_EINVAL, "rng is null");
        uint32_t mask = 0;
        if (results->inventory || results->inject_withdraw || results->consumed || results->loss || results->net_volume || results->pv) mask |= 1;
        if (results->decision_index) mask |= 2;
        ctx->outputs = mask;
        const bool verbose = getenv("LSMC_VERBOSE") != nullptr;   // development: host-side time of each stage
        auto now = [] { return std::chrono::steady_clock::now(); };
        auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        const auto t0 = now();
        set_problem(ctx, problem);
        // every allocation the passes below would make lazily: no cudaMalloc / cudaFree (device-wide synchronisation) once
        // the ranks of a group have started to wait for each other on the device
        ctx->fwd_params.alloc((size_t)(ctx->plan.P - 1));
        auto injected = [](const lsmc_rng* r) { return r->mode == LSMC_RNG_NORMALS || r->mode == LSMC_RNG_MT19937 || r->mode == LSMC_RNG_MT19937_CONTINUE; };
        if (injected(regression_rng) || injected(valuation_rng))
            ctx->normals_stage.alloc((size_t)ctx->plan.S * ctx->plan.F * (size_t)ctx->N);
        setup_done = true;
        if (after_setup && !after_setup()) throw Error(LSMC_ECANCELLED, "Another GPU of the group failed during set-up.");
        const auto t1 = now();
        simulate(ctx, 0, regression_rng);
        backward(ctx, progress, user);
        simulate(ctx, 1, valuation_rng);
        forward(ctx, progress, user);
        const auto t2 = now();
        get_results(ctx, results);
        if (verbose) fprintf(stderr, "lsmc_calculate: set_problem %.2f ms, simulate+backward+forward %.2f ms, get_results %.2f ms\n", ms(t0, t1), ms(t1, t2), ms(t2, now()));
    });
    if (rc != LSMC_OK && !setup_done && after_setup) { if (ctx->group) ctx->group->cancelled.store(1); after_setup(); }   // keep the group's barrier count whole
    return rc;
}

int lsmc_calculate(lsmc_ctx* ctx, const lsmc_problem* problem, const lsmc_rng* regression_rng, const lsmc_rng* valuation_rng,
                   lsmc_progress_cb progress, void* user, lsmc_results* results) {
    return calculate_impl(ctx, problem, regression_rng, valuation_rng, progress, user, results, nullptr);
}

int lsmc_host_alloc(size_t bytes, void** out) {
    if (!out) return LSMC_EINVAL;
    *out = nullptr;
    return guarded(nullptr, [&] { CUDA_CHECK(cudaHostAlloc(out, std::max<size_t>(bytes, 1), cudaHostAllocPortable)); });
}

void lsmc_host_free(void* p) { if (p) cudaFreeHost(p); }

int lsmc_intrinsic(const lsmc_problem* problem, double* npv, double* profile) {
    return guarded(nullptr, [&] {
        require(problem != nullptr && npv != nullptr, LSMC_EINVAL, "problem or npv is null");
        Plan plan;
        build_plan(*problem, plan);
        *npv = intrinsic_value(plan, profile);
    });
}

int lsmc_get_timings(const lsmc_ctx* ctx, lsmc_timings* out) {
    if (!ctx || !out) return LSMC_EINVAL;
    *out = ctx->tm;
    return LSMC_OK;
}

int lsmc_mt19937_normals(uint64_t seed, int antithetic, int64_t n_sims, int32_t n_steps, int32_t n_factors, double* out) {
    if (!out || n_sims < 1 || n_steps < 1 || n_factors < 1) return LSMC_EINVAL;
    Mt19937 g;
    g.seed((uint32_t)seed);
    mt19937_path_normals(g, antithetic != 0, n_sims, 0, n_sims, n_steps, n_factors, out);
    return LSMC_OK;
}

int lsmc_set_kernel_timing(lsmc_ctx* ctx, int on) {
    if (!ctx) return LSMC_EINVAL;
    ctx->time_kernels = on != 0;
    return LSMC_OK;
}

int lsmc_get_grid(const lsmc_ctx* ctx, int period, double* out, int32_t* count) {
    if (!ctx || !ctx->has_problem || period < 0 || period >= ctx->plan.P || !count) return LSMC_EINVAL;
    const int n = ctx->plan.grid_size(period);
    *count = n;
    if (out) std::memcpy(out, ctx->plan.grid_of(period), sizeof(double) * n);
    return LSMC_OK;
}

int lsmc_get_inventory_space(const lsmc_ctx* ctx, double* min_out, double* max_out) {
    if (!ctx || !ctx->has_problem || !min_out || !max_out) return LSMC_EINVAL;
    std::memcpy(min_out, ctx->plan.space_min.data(), sizeof(double) * ctx->plan.P);
    std::memcpy(max_out, ctx->plan.space_max.data(), sizeof(double) * ctx->plan.P);
    return LSMC_OK;
}

int lsmc_get_regression_coeffs(lsmc_ctx* ctx, int period, double* out) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] {
        require(ctx->backward_done && period >= 0 && period < ctx->plan.P - 1 && out, LSMC_EINVAL, "bad period or backward not run");
        const int Gn = ctx->plan.grid_size(period + 1), K = ctx->plan.K;
        std::vector<double> tmp((size_t)Gn * ctx->ldb);
        CUDA_CHECK(cudaMemcpy(tmp.data(), ctx->beta_all.p + (size_t)period * ctx->beta_stride(), tmp.size() * sizeof(double), cudaMemcpyDeviceToHost));
        for (int i = 0; i < Gn; i++) for (int k = 0; k < K; k++) out[(size_t)i * K + k] = tmp[(size_t)i * ctx->ldb + k];
    });
}

int lsmc_get_sim_panel(lsmc_ctx* ctx, int set, int factor, double* out) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] {
        require(ctx->has_problem && (set == 0 || set == 1) && ctx->sims_ready[set] && out, LSMC_EINVAL, "simulation set not available");
        const Plan& pl = ctx->plan;
        require(factor >= -1 && factor < pl.F, LSMC_EINVAL, "bad factor index");
        require(ctx->out_pitch == 0, LSMC_EINVAL, "lsmc_get_sim_panel is a single-context call");
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        if (factor < 0) { copy_panel(ctx, out, ctx->spot[set].p, pl.S); CUDA_CHECK(cudaStreamSynchronize(ctx->stream)); ctx->export_pending = false; return; }
        require(ctx->fac_set == set, LSMC_EINVAL, "the factors of this set were overwritten by the other simulation set");
        for (int j = 0; j < pl.S; j++)
            copy_panel(ctx, out + (size_t)j * ctx->N, ctx->fac.p + ((size_t)j * pl.F + factor) * ctx->Npad, 1);
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream)); ctx->export_pending = false;
    });
}

int lsmc_debug_set_regression_coeffs(lsmc_ctx* ctx, int period, const double* coeffs) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] {
        require(ctx->has_problem && period >= 0 && period < ctx->plan.P - 1 && coeffs, LSMC_EINVAL, "bad period");
        if (ctx->beta_override.empty()) ctx->beta_override.assign((size_t)(ctx->plan.P - 1) * ctx->beta_stride(), 0.0);
        const int Gn = ctx->plan.grid_size(period + 1), K = ctx->plan.K;
        double* dst = ctx->beta_override.data() + (size_t)period * ctx->beta_stride();
        std::fill(dst, dst + ctx->beta_stride(), 0.0);
        for (int i = 0; i < Gn; i++) for (int k = 0; k < K; k++) dst[(size_t)i * ctx->ldb + k] = coeffs[(size_t)i * K + k];
        ctx->beta_override_set[period] = 1;
    });
}

int lsmc_debug_keep_values(lsmc_ctx* ctx, int keep) {
    if (!ctx) return LSMC_EINVAL;
    ctx->keep_values = keep != 0;
    return LSMC_OK;
}

int lsmc_debug_get_values(lsmc_ctx* ctx, int period, double* out) {
    if (!ctx) return LSMC_EINVAL;
    return guarded(ctx, [&] {
        require(ctx->backward_done && ctx->keep_values && ctx->Vall.p && period >= 0 && period < ctx->plan.P && out, LSMC_EINVAL, "values were not kept");
        copy_panel(ctx, out, ctx->Vall.p + (size_t)period * ctx->plan.max_grid * ctx->Npad, ctx->plan.grid_size(period));
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream)); ctx->export_pending = false;
    });
}

}  // extern "C"
