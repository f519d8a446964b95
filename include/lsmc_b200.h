/*
 * lsmc_b200.h — C ABI of the B200-native Least-Squares Monte Carlo storage valuation path.
 *
 * This is the drop-in boundary for the hot path of Cmdty.Storage's
 *   LsmcStorageValuation.Calculate<T>(LsmcValuationParameters<T>)
 *     (reference: src/Cmdty.Storage/LsmcValuation/LsmcStorageValuation.cs:55-617)
 * i.e. everything between "inputs validated" (LSV.cs:95) and "results object constructed"
 * (LSV.cs:614): spot simulation (LVP.cs:160-178), backward induction (LSV.cs:104-341),
 * forward simulation (LSV.cs:344-579), NPV / deltas / profile / trigger prices (LSV.cs:482-561,584-595).
 *
 * The host (C# via P/Invoke, or Python via ctypes) flattens LsmcValuationParameters<T> into the plain
 * arrays below; no time-period, delegate or managed type crosses this boundary.  All arrays are
 * caller-owned; the library copies what it needs during the call and keeps no host pointer.
 * All functions return an lsmc_status; never throw.  A context is single-caller (not thread safe),
 * distinct contexts are independent (reference threading: ExcelCalcWrapper.cs:57-67).
 *
 * Period index convention: p = 0 .. P-1 enumerates `periodsForResultsTimeSeries`
 * (LSV.cs:154) = startActiveStorage .. storage.EndPeriod inclusive.  Decision periods are p < P-1.
 */
#ifndef LSMC_B200_H
#define LSMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Status codes and the .NET exception each one maps to. */
typedef enum lsmc_status {
    LSMC_OK = 0,
    LSMC_EINVAL = 1,        /* ArgumentException (LSV.cs:62,92,95; SH.cs:113-116) */
    LSMC_ECONSTRAINT = 2,   /* InventoryConstraintsCannotBeFulfilledException (SH.cs:102,131,155; CmdtyStorage.cs:88-94) */
    LSMC_ECANCELLED = 3,    /* OperationCanceledException (LSV.cs:339,490) */
    LSMC_ECUDA = 4,         /* ApplicationException: CUDA / NCCL failure */
    LSMC_EUNSUPPORTED = 5   /* NotSupportedException: storage/basis form without a device descriptor */
} lsmc_status;

/* Inject/withdraw constraint kinds (src/Cmdty.Storage/InjectWithdrawConstraints/). */
enum {
    LSMC_CONSTRAINT_CONSTANT = 0,          /* ConstantInjectWithdrawConstraint.cs:45 */
    LSMC_CONSTRAINT_PIECEWISE_LINEAR = 1,  /* PiecewiseLinearInjectWithdrawConstraint.cs:67 */
    LSMC_CONSTRAINT_STEP = 2,              /* StepInjectWithdrawConstraint.cs:72 */
    LSMC_CONSTRAINT_POLYNOMIAL = 3         /* PolynomialInjectWithdrawConstraint.cs:80: rates are polynomials in the inventory.  The host
                                              fits them (the C# shim passes MathNet's own Fit.Polynomial coefficients, :66-73):
                                              pillar_min_rate / pillar_max_rate hold the COEFFICIENTS c_0 .. c_{n-1} (ascending powers,
                                              evaluated by Horner's rule like Polynomial.Evaluate), pillar_inventory the fitted pillars'
                                              inventories (ascending; only the first and last are used, as the inventory range) */
};

/* Normal-draw source for one simulation set. */
enum {
    LSMC_RNG_PHILOX = 0,    /* device Philox4x32-10 + Box-Muller */
    LSMC_RNG_NORMALS = 1,   /* caller-injected standard normals (parity mode) */
    LSMC_RNG_PANELS = 2,    /* caller-injected spot/factor panels (e.g. from Cmdty.Core's own simulator) */
    LSMC_RNG_MT19937 = 3,   /* the reference's seeded stream (LsmcValuationParameters.cs:181-191: MersenneTwisterGenerator(seed,
                               antithetic)): MathNet MT19937 + Normal.Samples, one generator call per path, generated serially on
                               the host (csrc/mt19937.h) and injected; reproduces the reference's seeded golden values */
    LSMC_RNG_MT19937_CONTINUE = 4   /* the same generator continues after the context's previous MT19937 simulation
                                       (LVP.cs:186-187: valuationSimSeed == null) */
};

/*
 * Flattened LsmcValuationParameters<T> (LVP.cs:42-56) + ICmdtyStorage<T> (ICmdtyStorage.cs:32) +
 * MultiFactorParameters<T> (Cmdty.Core).  Per-period arrays have P = n_periods entries.
 */
typedef struct lsmc_problem {
    int32_t struct_size;              /* = sizeof(lsmc_problem); ABI guard */
    int32_t n_periods;                /* P (LSV.cs:104,154) */
    int32_t n_factors;                /* F, 1..LSMC_MAX_FACTORS */
    int32_t n_basis;                  /* K, 1..LSMC_MAX_BASIS */
    int64_t n_sims;                   /* N: paths owned by THIS context (the local shard under multi-GPU) */
    int64_t n_sims_global;            /* total paths over all ranks; 0 => n_sims */
    int64_t sim_offset;               /* global index of this context's first path (Philox counters) */
    int32_t extra_decisions;          /* LVP.cs:56 */
    int32_t first_period_is_current;  /* 1 when period 0 == CurrentPeriod: deterministic price, not simulated (LSV.cs:171,226,382,410; LVP.cs:173) */
    int32_t discount_deltas;          /* LVP.cs:55 */
    int32_t must_be_empty_at_end;     /* storage.MustBeEmptyAtEnd (CmdtyStorage.cs:52) */
    double  numerical_tolerance;      /* LVP.cs:49 */
    double  start_inventory;          /* LVP.cs:43 */
    double  grid_spacing;             /* FixedSpacingStateSpaceGridCalc.Spacing (FixedSpacingStateSpaceGridCalc.cs:36) */
    double  terminal_coef[4];         /* TerminalStorageNpv(price, inv) = c0 + c1*price + c2*inv + c3*(price*inv); used when !must_be_empty_at_end */

    /* price model */
    const double* sim_time;           /* [P] Act/365 years from CurrentPeriod.Start to the start of period p (LVP.cs:172-176) */
    const double* mean_reversion;     /* [F] */
    const double* factor_vol;         /* [P][F] vol of factor f for delivery period p */
    const double* factor_corr;        /* [F][F] */
    const double* forward;            /* [P] ForwardCurve[period p] */

    /* regression basis: column k = spot^spot_power[k] * prod_f factor_f^factor_power[k][f] (BasisFunctions.cs:92-132) */
    const int32_t* spot_power;        /* [K] */
    const int32_t* factor_power;      /* [K][F] */

    /* storage, per period */
    const double* min_inventory;      /* [P] storage.MinInventory(period) */
    const double* max_inventory;      /* [P] storage.MaxInventory(period) (0 at the end period when must-be-empty, CmdtyStorage.cs:432-438) */
    const int32_t* constraint_id;     /* [P] index into the constraint tables */
    const double* inject_cost;        /* [P] per-unit injection cost */
    const double* inject_cost_df;     /* [P] discount factor to the injection-cost cash-flow date */
    const double* withdraw_cost;      /* [P] */
    const double* withdraw_cost_df;   /* [P] */
    const double* consumed_inject;    /* [P] fraction of commodity consumed on injection */
    const double* consumed_withdraw;  /* [P] */
    const double* inventory_loss;     /* [P] fractional inventory loss per period */
    const double* inventory_cost;     /* [P] per-unit inventory cost */
    const double* inventory_cost_df;  /* [P] */
    const double* settle_df;          /* [P] discount factor to SettleDateRule(period) (LSV.cs:222-223) */

    /* constraint tables */
    int32_t n_constraints;            /* C */
    int32_t n_pillars;                /* total pillar rows */
    const int32_t* constraint_kind;   /* [C] LSMC_CONSTRAINT_* */
    const int32_t* constraint_first;  /* [C] first pillar row */
    const int32_t* constraint_count;  /* [C] pillar rows (1 for constant) */
    const double* pillar_inventory;   /* [n_pillars] ascending within one constraint */
    const double* pillar_min_rate;    /* [n_pillars] MinInjectWithdrawRate (<= 0 is withdrawal) */
    const double* pillar_max_rate;    /* [n_pillars] */
} lsmc_problem;

#define LSMC_MAX_FACTORS 8
#define LSMC_MAX_BASIS 64
#define LSMC_NUM_TRIGGER_VOLUMES 10   /* LSV.cs:367 */
#define LSMC_MAX_POLYNOMIAL_PILLARS 16 /* polynomial constraints: at most degree 15 */

typedef struct lsmc_rng {
    int32_t mode;             /* LSMC_RNG_* */
    int32_t antithetic;       /* Philox mode: global paths 2m and 2m+1 share draw m with opposite signs */
    uint64_t seed;            /* Philox key; MT19937: the low 32 bits are MathNet's int seed */
    const double* normals;    /* LSMC_RNG_NORMALS: [S][F][N] standard normals, S = simulated periods */
    const double* spot;       /* LSMC_RNG_PANELS: [S][N] */
    const double* factors;    /* LSMC_RNG_PANELS: [S][F][N] */
} lsmc_rng;

/* Sizes the caller needs to allocate result buffers. */
typedef struct lsmc_layout {
    int32_t n_periods;        /* P */
    int32_t n_sim_periods;    /* S = P - first_period_is_current */
    int32_t max_grid;         /* largest inventory grid over all periods */
    int32_t max_decisions;    /* 2*extra_decisions + 3 */
    int64_t n_sims;
} lsmc_layout;

/*
 * Results (LsmcStorageValuationResults.cs:34-74).  Pointers may be NULL to skip that output.
 * Panels are row = period, column = sim, exactly like Panel<T,double> (utils.py:315-320).
 */
typedef struct lsmc_results {
    int32_t struct_size;
    int32_t reserved;
    double npv;                /* mean of pv_by_sim (LSV.cs:584) */
    double npv_std_error;      /* standard error of npv (antithetic pair means when antithetic) */
    double backward_npv;       /* mean of the backward-induction value at the start inventory (LSV.cs:589) */
    double* deltas;            /* [P] (LSV.cs:360,486-487; last entry 0) */
    double* profile;           /* [P][6] inventory, inject_withdraw, consumed, loss, net_volume, period_pv (StorageProfile.cs:28) */
    double* trigger_summary;   /* [P-1][4] max inject volume, max inject price, max withdraw volume, max withdraw price; NaN = null (TriggerPrices.cs:30-33) */
    double* trigger_inject;    /* [P-1][10][2] (volume, price), NaN rows when absent (LSV.cs:517-523) */
    double* trigger_withdraw;  /* [P-1][10][2] (LSV.cs:545-551) */
    double* pv_by_sim;         /* [N] */
    double* spot_regress;      /* [S][N] RegressionSpotPriceSim */
    double* spot_valuation;    /* [S][N] ValuationSpotPriceSim */
    double* inventory;         /* [P][N] InventoryBySim */
    double* inject_withdraw;   /* [P][N] */
    double* consumed;          /* [P][N] */
    double* loss;              /* [P][N] */
    double* net_volume;        /* [P][N] */
    double* pv;                /* [P][N] PvByPeriodAndSim */
    int8_t* decision_index;    /* [P-1][N] index of the chosen decision in the bang-bang set (parity checks) */
} lsmc_results;

/* Device-side phase timers, the CUDA-event equivalent of Stopwatches.cs:35-40. */
typedef struct lsmc_timings {
    double regression_sim_ms;
    double backward_ms;
    double valuation_sim_ms;
    double forward_ms;
    double total_ms;
    double backward_sweep_kernel_ms;   /* sum over periods of the fused decision+normal-equation kernel; 0 unless
                                          lsmc_set_kernel_timing(ctx, 1) was called before lsmc_backward */
    double backward_solve_kernel_ms;   /* sum over periods of reduce+solve */
    int64_t backward_sweep_launches;
    int64_t kernel_launches;           /* all kernels launched since lsmc_set_problem */
    double sweep_bytes;                /* algorithmic HBM bytes moved by the sweep kernels */
    double span_ms;                    /* one CUDA-event interval from the start of the regression simulation to the end of the
                                          forward pass on the context's stream (phase gaps included); 0 until all four phases ran */
} lsmc_timings;

typedef struct lsmc_ctx lsmc_ctx;

/* Progress / cancellation: called on the calling thread once per time step with the reference's
 * progress value (LSV.cs:337-338,488-489,604); return non-zero to cancel (=> LSMC_ECANCELLED). */
typedef int (*lsmc_progress_cb)(void* user, double progress);

/* Context: owns one CUDA device, its stream and all device memory until destroyed. */
int  lsmc_create(int device, lsmc_ctx** out);
void lsmc_destroy(lsmc_ctx* ctx);
const char* lsmc_last_error(const lsmc_ctx* ctx);   /* ctx may be NULL: last error of a failed lsmc_create */
const char* lsmc_version(void);

/* Multi-GPU, one process per GPU: paths are sharded, only the normal-equation partials and the forward sums are
 * exchanged.  lsmc_comm_init creates the NCCL communicator (NCCL is loaded with dlopen on first use) and then maps every
 * rank's exchange buffer into every other rank (cudaIpc): the per-period exchange is fused into the kernel that reduces
 * and solves the normal equations and goes over NVLink peer memory; NCCL collectives remain as the fallback when the
 * buffers cannot be mapped. */
int lsmc_comm_unique_id(void* out128);                                  /* ncclGetUniqueId */
int lsmc_comm_init(lsmc_ctx* ctx, int rank, int world, const void* unique_id128);

/* Multi-GPU inside ONE process (the shape of the reference: Calculate<T> is one call on one thread):
 * lsmc_comm_init_local joins n contexts (one per GPU; the same GPU may appear twice, for tests) into a group whose
 * exchanges go through peer memory only (no NCCL); lsmc_calculate_multi takes the WHOLE problem (n_sims = all paths),
 * splits the paths over the contexts (one internal host thread per GPU, rank 0 on the calling thread: progress callbacks
 * arrive there) and fills `results` as lsmc_calculate would: summary from rank 0, per-sim outputs [rows][n_sims] with
 * every rank writing its own block of columns.  Injected normals / panels are sliced per rank. */
int lsmc_comm_init_local(lsmc_ctx** ctxs, int n);
int lsmc_calculate_multi(lsmc_ctx** ctxs, int n, const lsmc_problem* problem,
                         const lsmc_rng* regression_rng, const lsmc_rng* valuation_rng,
                         lsmc_progress_cb progress, void* progress_user, lsmc_results* results);

/* One-call entry point: what the body of Calculate<T> becomes. */
int lsmc_calculate(lsmc_ctx* ctx, const lsmc_problem* problem,
                   const lsmc_rng* regression_rng, const lsmc_rng* valuation_rng,
                   lsmc_progress_cb progress, void* progress_user, lsmc_results* results);

/* Staged entry points (same work, separately callable; used by the benchmark and the tests). */
int lsmc_set_problem(lsmc_ctx* ctx, const lsmc_problem* problem);       /* host set-up (SH.cs:39-107, grids, decision tables) + H2D */
int lsmc_get_layout(const lsmc_ctx* ctx, lsmc_layout* out);
int lsmc_request_outputs(lsmc_ctx* ctx, uint32_t mask);                 /* before lsmc_set_problem: bit0 per-sim [P][N] panels, bit1 decision_index */
int lsmc_simulate(lsmc_ctx* ctx, int set /*0 regression, 1 valuation*/, const lsmc_rng* rng);
int lsmc_backward(lsmc_ctx* ctx, lsmc_progress_cb progress, void* progress_user);
int lsmc_forward(lsmc_ctx* ctx, lsmc_progress_cb progress, void* progress_user);
int lsmc_get_results(lsmc_ctx* ctx, lsmc_results* results);             /* D2H of whatever is non-NULL */
/* The standard normals LSMC_RNG_MT19937 injects for a run of n_sims paths x n_steps x n_factors, [n_steps][n_factors][n_sims]
 * (host only, no GPU needed): what Cmdty.Core's MultiFactorSpotPriceSimulator draws from MersenneTwisterGenerator(seed, antithetic). */
int lsmc_mt19937_normals(uint64_t seed, int antithetic, int64_t n_sims, int32_t n_steps, int32_t n_factors, double* out);
int lsmc_get_timings(const lsmc_ctx* ctx, lsmc_timings* out);
/* Per-launch CUDA events around every sweep kernel of the following lsmc_backward calls (fills backward_sweep_kernel_ms;
 * the reference's Stopwatches.cs has no finer counterpart).  Off by default: an event between two kernels serialises them,
 * and the kernels of the backward loop otherwise overlap their set-up with the predecessor's drain (programmatic
 * dependent launch), about 5 us per period. */
int lsmc_set_kernel_timing(lsmc_ctx* ctx, int on);

/* Page-locked host memory for result buffers.  lsmc_get_results / lsmc_calculate write per-sim panels ([P][N] doubles each:
 * 23 GB for the eight panels of a 1M-path daily year, LSV.cs:351-356,602-616) with the copy engine straight into a
 * destination that is page-locked (these helpers, cudaHostAlloc or cudaHostRegister) — all panels in flight, one
 * synchronisation; a pageable destination is served through two 32 MB page-locked staging chunks, the copy of one chunk
 * overlapping the host memcpy of the other. */
int  lsmc_host_alloc(size_t bytes, void** out);
void lsmc_host_free(void* p);

/* Intrinsic valuation on the forward curve, the deterministic model the reference's Python entry points run beside the
 * LSMC (multi_factor.py:406 -> IntrinsicStorageValuation.cs:120-322): backward induction over the same inventory grids with
 * linear interpolation of the value function, then the optimal schedule from the start inventory.  Host only (O(P * G));
 * needs no context and no GPU.  profile: [P][6] StorageProfile rows (inventory AFTER the period's decision) or NULL.
 * Only the storage / curve fields of the problem are read (n_sims >= 1, any basis / factor model).  Errors: lsmc_last_error(NULL). */
int lsmc_intrinsic(const lsmc_problem* problem, double* npv, double* profile /*[P][6] or NULL*/);

/* Introspection used by hosts that want RegressionCoefficients / InventoryGrids (commented out in
 * LsmcStorageValuationResults.cs:51-52) and by the parity tests. */
int lsmc_get_grid(const lsmc_ctx* ctx, int period, double* out /*[max_grid]*/, int32_t* count);
int lsmc_get_inventory_space(const lsmc_ctx* ctx, double* min_out /*[P]*/, double* max_out /*[P]*/);
int lsmc_get_regression_coeffs(lsmc_ctx* ctx, int period, double* out /*[grid(period+1)][K]*/);
int lsmc_get_sim_panel(lsmc_ctx* ctx, int set, int factor /*-1 = spot*/, double* out /*[S][N]*/);

/*
 * Allocation-free NPV statistics.  This entry point has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It is the NPV / standard-error reduction that closes lsmc_forward (mean of pv_by_sim,
 * LSV.cs:584, and the standard error over antithetic pair means), with the block partials in a file-scope __device__
 * array of the library (32 KB per GPU that loads the library, held until exit) instead of cudaMalloc'ed scratch.
 *
 *   d_pv_by_sim       [n_sims] DEVICE pointer (e.g. the caller's own copy of lsmc_results.pv_by_sim)
 *   antithetic_pairs  non-zero: paths 2m and 2m+1 are antithetic partners, the standard error is taken over pair means
 *                     (only when n_sims is even, as lsmc_forward does)
 *   d_out4            [4] DEVICE pointer: npv, npv_std_error, sum of pv, sum of squared deviations
 *   cuda_stream       cudaStream_t of the CURRENT device (NULL = default stream); the call only enqueues three kernels
 *   err, err_len      optional caller buffer for the error text
 *
 * Calls no cudaMalloc / cudaFree / allocating library routine.  n_sims above LSMC_NOALLOC_MAX_SIMS is rejected with
 * LSMC_EUNSUPPORTED.  All calls on one GPU share the static array (it keeps the last call's contents; nothing in it has to
 * start at zero): calls that overlap on the GPU race — serialise them on one stream.  The first call on a device looks the
 * array's device address up (cudaGetSymbolAddress); make that call outside stream capture.
 */
#define LSMC_NOALLOC_MAX_SIMS (1 << 24)   /* 16,777,216 paths */
int lsmc_npv_stats_noalloc(const double* d_pv_by_sim, int64_t n_sims, int antithetic_pairs, double* d_out4,
                           void* cuda_stream, char* err, size_t err_len);

/* Test hooks (tests/ only). */
int lsmc_debug_set_regression_coeffs(lsmc_ctx* ctx, int period, const double* coeffs /*[grid(period+1)][K]*/);
int lsmc_debug_keep_values(lsmc_ctx* ctx, int keep);                    /* keep every period's value surface */
int lsmc_debug_get_values(lsmc_ctx* ctx, int period, double* out /*[grid(period)][N]*/);

#ifdef __cplusplus
}
#endif
#endif /* LSMC_B200_H */
