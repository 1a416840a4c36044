/*
 * mc_b200.h -- C ABI of the B200-native Monte Carlo path simulator.
 *
 * This is the drop-in boundary for the one hot path of raasawin/montecarlo:
 * the routine `garch_monte_carlo(df, n_sims, n_days, current_price,
 * progress_callback=None)` (reference app.py:586-657, sole call site
 * app.py:1240).  The reference has no FFI of its own -- its boundary is that
 * plain Python function -- so the entry points below are what a ctypes binding
 * for it needs: one call that takes the bootstrap pool and the scalar model
 * parameters and returns every reduction the function's result dict holds
 * (app.py:647-657), plus a staged variant that lets the host interleave one
 * small collective per stage when paths are sharded over several GPUs.
 *
 * Conventions: plain C, no C++ types or exceptions cross the boundary.  Every
 * function returns 0 on success, a negative mc_status otherwise;
 * mc_last_error(h) gives the message.  All `d_` pointers are DEVICE pointers
 * owned by the caller (PyTorch tensors' data_ptr()), `h_` pointers are host
 * memory.  Kernels are enqueued on the caller's CUDA stream (passed as void*,
 * a cudaStream_t); only mc_fetch_result / mc_run synchronise.  A handle is
 * bound to one device and guarded by a mutex: concurrent calls on different
 * handles are safe (the Streamlit app runs sessions in threads, app.py:1392).
 */
#ifndef MC_B200_H
#define MC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MC_ABI_VERSION 3u
#define MC_MAX_PCTS 16u       /* terminal percentiles per call (reference uses 9, app.py:639) */
#define MC_MAX_PATH_PCTS 8u   /* fan-chart percentiles (reference uses 5, app.py:644) */
#define MC_MAX_ASSETS 32u

typedef enum mc_status {
    MC_OK = 0,
    MC_ERR_INVALID = -1,     /* bad argument / unsupported combination */
    MC_ERR_CUDA = -2,        /* a CUDA runtime call failed */
    MC_ERR_VERSION = -3,     /* mc_params.struct_size / version mismatch */
    MC_ERR_STATE = -4        /* stage functions called out of order */
} mc_status;

/* where the per-step shock comes from */
typedef enum mc_shock {
    MC_SHOCK_BOOTSTRAP = 0,  /* Philox word -> index into the pool: replaces np.random.choice(std_returns) app.py:620 */
    MC_SHOCK_NORMAL = 1,     /* Philox -> Box-Muller N(0,1) in registers (north_star extension) */
    MC_SHOCK_INJECTED = 2    /* read z[t, path] from d_shocks (parity testing) */
} mc_shock;

/* how the per-step variance evolves */
typedef enum mc_var_model {
    MC_VAR_GARCH = 0,        /* var' = omega + alpha r^2 + beta var   app.py:632-633 */
    MC_VAR_CONST = 1         /* constant variance var0 (GBM) */
} mc_var_model;

typedef enum mc_dtype { MC_F32 = 0, MC_F64 = 1 } mc_dtype;

enum {
    MC_FLAG_STORE_PATHS = 1u,   /* write paths[(T+1), ld] time-major like app.py:612-613 */
    MC_FLAG_FAN_CHART = 2u,     /* per-step percentiles (app.py:643-645); needs STORE_PATHS */
    MC_FLAG_NO_TERMINAL_SELECT = 4u, /* skip percentiles / tail (moments + histogram only) */
    MC_FLAG_PATHS_VOLATILE = 8u, /* staged runs: d_paths is shared with other runs and may be overwritten between
                                   stages (fan chart by regeneration).  From the second narrowing pass on, the
                                   handle simulates the run again by itself, and only if the values it kept from
                                   pass 1 do not cover what the pass needs; mc_stage_reads_paths then reports 1
                                   only for the first pass. */
    MC_FLAG_FULL_RES_NORMALS = 32u, /* fp32 normal shocks from full 32-bit uniforms, four per Philox call (|z| <= 6.66,
                                   directions through MUFU.SIN/COS) instead of the default eight per call (20-bit radius:
                                   |z| <= 5.27, P(|z| > 5) 27 % low; 4096 tabulated directions).  ~25 % slower; for callers
                                   who look at single-step tails.  Single-asset fp32 kernels; fp64 always draws this way. */
    MC_FLAG_TERMINAL_GIVEN = 16u /* d_terminal already holds this rank's final prices (app.py:636): no simulation, only
                                   the reductions of app.py:638-640 and 652-656.  What mc_run itself falls back to when
                                   the fused closing reports mc_result.status = 1; the same call serves a caller who
                                   produced the terminal values elsewhere. */
};

/*
 * One simulation request.  Everything the reference hard-codes or derives
 * before its loop (app.py:601-609) arrives here as a number.
 */
typedef struct mc_params {
    uint32_t struct_size;       /* sizeof(mc_params) */
    uint32_t version;           /* MC_ABI_VERSION */
    uint64_t n_paths;           /* paths simulated by THIS call (this rank's shard) */
    uint64_t path_offset;       /* global id of local path 0: Philox counter = global id */
    uint64_t n_paths_global;    /* paths over all ranks (order statistics are global) */
    uint64_t paths_ld;          /* row stride (elements) of d_paths, >= n_paths; 0 = n_paths */
    uint32_t n_steps;           /* n_days */
    uint32_t n_assets;          /* 1, or d <= MC_MAX_ASSETS for the portfolio mode */
    int32_t shock;              /* mc_shock */
    int32_t var_model;          /* mc_var_model */
    int32_t dtype;              /* mc_dtype of terminal / paths / pool / shocks buffers */
    uint32_t flags;
    uint64_t seed;              /* Philox key */
    double s0;                  /* current_price            app.py:613 */
    double var0;                /* initial per-step variance app.py:606 (GBM: sigma^2 dt) */
    double omega, alpha, beta;  /* app.py:601-603 */
    double drift;               /* log-space drift per step (reference: 0; GBM: (mu - sigma^2/2) dt) */
    uint32_t pool_len;          /* H = len(std_returns)      app.py:609 */
    uint32_t hist_bins;         /* 0 = no histogram */
    double hist_lo, hist_hi;    /* np.histogram range; lo >= hi means (min, max) of the data */
    uint32_t n_pcts;            /* terminal percentiles, ascending, in percent */
    uint32_t n_path_pcts;
    double pcts[MC_MAX_PCTS];
    double path_pcts[MC_MAX_PATH_PCTS];
    double tail_pct;            /* CVaR level in percent (reference: 5, app.py:656); <0 = none */
    uint32_t n_ranks;           /* ranks sharing n_paths_global (0 or 1 = single GPU); sizes the per-rank pilot sample */
    uint32_t pass_group;        /* 0 / 1: none.  g > 1 (fan chart by regeneration): g handles of this GPU are virtual ranks of one
                                   job and write their narrowing-pass outputs into ONE buffer (mc_stage_adopt); a gathered pass
                                   stage then holds n_ranks / g buffers.  Must divide n_ranks. */
} mc_params;

/* Portfolio description (n_assets > 1); host arrays, copied at launch. */
typedef struct mc_portfolio {
    const double* h_chol;       /* [d*d] row-major lower-triangular Cholesky factor of the correlation */
    const double* h_sigma_step; /* [d] sigma_i * sqrt(dt) */
    const double* h_drift_step; /* [d] (mu_i - sigma_i^2/2) dt */
    const double* h_s0;         /* [d] initial prices */
    const double* h_weights;    /* [d] portfolio weights */
} mc_portfolio;

/* Everything scalar the reference's result dict holds (app.py:647-657). */
typedef struct mc_result {
    uint64_t n;                 /* paths reduced (global) */
    uint64_t n_above;           /* #(final > s0)                    -> prob_profit  app.py:654 */
    double mean;                /* mean(final)                      -> expected_return app.py:652 */
    double std;                 /* std(final), ddof=0               -> expected_vol app.py:653 */
    double min, max;
    double pct[MC_MAX_PCTS];    /* np.percentile(final, p), method='linear'  app.py:638-640 */
    double pct_lo[MC_MAX_PCTS]; /* the two order statistics each percentile interpolates */
    double pct_hi[MC_MAX_PCTS];
    uint64_t tail_count;        /* #(final <= percentile(tail_pct)) */
    double tail_sum;            /* their sum                        -> cvar_95 app.py:656 */
    double hist_lo, hist_hi;    /* range actually used */
    float sim_ms;               /* device time of the simulation kernel(s), CUDA events */
    float total_ms;             /* device time of the whole call */
    uint32_t launches;          /* kernels launched by this call */
    uint32_t status;            /* 0: complete.  1 (mc_run_sharded only; mc_run handles it itself): the key window of the
                                   fused closing missed or a candidate list overflowed -- identically on every rank --
                                   and the reductions must be redone with mc_stage_* and MC_FLAG_TERMINAL_GIVEN */
} mc_result;

typedef struct mc_handle mc_handle;

int mc_abi_version(void);
int mc_create(int device_ordinal, mc_handle** out);
int mc_destroy(mc_handle* h);
const char* mc_last_error(const mc_handle* h);   /* h may be NULL: last create() error */

/*
 * Whole job on one GPU (replaces the body of garch_monte_carlo, app.py:612-656).
 *   d_pool     : [pool_len] std_returns (dtype), MC_SHOCK_BOOTSTRAP only
 *   d_shocks   : [n_steps (* n_assets), ld] injected shocks, MC_SHOCK_INJECTED only
 *   d_paths    : [(n_steps+1), paths_ld] or NULL          (MC_FLAG_STORE_PATHS)
 *   d_terminal : [n_paths] final prices (always written)
 *   h_result   : scalars;  h_hist: uint64[hist_bins] or NULL;
 *   h_fan      : double[n_path_pcts * (n_steps+1)] or NULL (MC_FLAG_FAN_CHART)
 * Synchronises `stream` before returning.  A request with both a fan chart and terminal percentiles runs the (one-row)
 * terminal selection on a stream the handle owns, forked from and joined back into `stream` by events: all of the
 * call's work is ordered after what `stream` held before the call and complete when the call returns.
 */
int mc_run(mc_handle* h, const mc_params* p, const mc_portfolio* pf,
           const void* d_pool, const void* d_shocks, void* d_paths, void* d_terminal,
           mc_result* h_result, uint64_t* h_hist, double* h_fan, void* stream);

/*
 * Staged variant for path-sharded multi-GPU runs.  Each stage writes
 * a packed buffer of 8-byte words; the host all-gathers it over ranks (NCCL) and
 * hands the gathered [n_ranks][words] buffer to the next stage.  With n_ranks=1
 * the "gathered" buffer is the stage buffer itself.  Stages: local moments + pilot
 * sample -> per-bracket counts/histograms (normally two narrowing passes; a pass
 * count that depends on the data is agreed on through a device flag every rank
 * computes identically) -> tail sums + fixed-bin histogram.
 *
 *   mc_stage_begin(...)                 simulate + local moments + pilot sample
 *   loop: words = mc_stage_words(h)     size of the buffer the next call writes/expects
 *         mc_stage_next(h, d_gathered, n_ranks, d_out, &done)
 *   mc_fetch_result(...)                D2H of the merged result, synchronises
 */
int mc_stage_begin(mc_handle* h, const mc_params* p, const mc_portfolio* pf,
                   const void* d_pool, const void* d_shocks, void* d_paths, void* d_terminal,
                   void* d_stage_out, uint64_t stage_capacity_words, void* stream);
int64_t mc_stage_words(const mc_handle* h);
int64_t mc_stage_max_words(const mc_params* p);
int mc_stage_next(mc_handle* h, const void* d_gathered, uint32_t n_ranks,
                  void* d_stage_out, int* done, void* stream);
int mc_fetch_result(mc_handle* h, mc_result* h_result, uint64_t* h_hist, double* h_fan,
                    void* stream);

/*
 * Path-sharded runs WITHOUT host round trips (runs without a fan chart).  The reductions of
 * app.py:636-656 need two exchanges between the ranks -- each rank's 32 KB key histogram of its terminal values, then
 * the values of the few histogram bins that hold the wanted order statistics plus moments / tail sums / fixed-bin
 * counts -- and the kernels carry them out themselves: every rank owns an exchange buffer, maps its peers' buffers
 * (CUDA IPC over NVLink), stores its contribution straight into every peer's buffer while it is being produced and
 * raises a flag there; the consuming kernel spins on its own copy of the flags.  No NCCL launch, no stage call.
 *
 *   mc_peer_export(h, n_ranks, ipc_handle[64], &local_ptr, &bytes)   allocate the buffer, name it
 *   (the host all-gathers the 64-byte handles: torch.distributed, MPI, files ...)
 *   mc_peer_attach(h, rank, n_ranks, all_handles, NULL)              map the peers (same process: pass the
 *                                                                    peers' local_ptr values instead of handles)
 *   mc_run_sharded(h, params, ...)        enqueue this rank's share; every rank calls it, in the same order
 *   mc_fetch_result(h, &result, hist, NULL, stream)                   synchronises; identical result on every rank
 *
 * mc_run_sharded needs params->n_ranks == n_ranks of the attachment and a request the fused closing accepts (no
 * fan chart, an explicit histogram range, <= 16384 bins); otherwise MC_ERR_INVALID and the staged protocol above
 * applies.  A peer that does not arrive within 20 s fails the run (MC_ERR_STATE from mc_fetch_result).
 *
 * With params->n_ranks <= 1 on a handle without peers, mc_run_sharded + mc_fetch_result are the two halves of mc_run
 * for the same requests: enqueue only, then wait and read.  A caller that alternates between two handles (each with
 * its own stream) keeps the GPU busy while a result travels; mc_result.status != 0 means the key window missed and
 * the reductions must be redone by mc_run with MC_FLAG_TERMINAL_GIVEN on the terminal values already there.
 */
int mc_peer_export(mc_handle* h, uint32_t n_ranks, void* out_ipc_handle /* 64 bytes, may be NULL */,
                   void** out_local_ptr, uint64_t* out_bytes);
int mc_peer_attach(mc_handle* h, uint32_t rank, uint32_t n_ranks, const void* ipc_handles /* n_ranks x 64 B or NULL */,
                   void* const* local_ptrs /* n_ranks device pointers (same process) or NULL */);
int mc_run_sharded(mc_handle* h, const mc_params* p, const mc_portfolio* pf,
                   const void* d_pool, const void* d_shocks, void* d_paths, void* d_terminal, void* stream);

/*
 * Fan chart without keeping the paths (app.py:643-645 at any n_sims).  A path range that
 * does not fit in HBM is cut into chunks that take turns in ONE [n_steps+1, chunk] buffer; each chunk runs
 * as a virtual rank of the staged protocol (one handle per chunk, n_ranks = ranks x chunks).  Paths are
 * a pure function of (seed, global path id), so before a stage that reads the rows again the chunk is
 * simply simulated again into the shared buffer:
 *
 *   if (mc_stage_reads_paths(h) == 1) mc_stage_resimulate(h, stream);
 *   mc_stage_next(h, ...);
 *
 * mc_stage_reads_paths: 1 if the next mc_stage_next of the run in flight streams the stored paths,
 * 0 if not, negative on error (it may wait for the device's "rows undecided" counter).
 */
int mc_stage_reads_paths(mc_handle* h);
int mc_stage_resimulate(mc_handle* h, void* stream);
/*
 * The chunks of one GPU all consume the same gathered buffer, and what a handle derives from it -- brackets, narrowed
 * intervals, the count of undecided rows -- does not depend on the handle.  One chunk (the leader) calls mc_stage_next;
 * the others call mc_stage_adopt with the same arguments plus the leader: they copy its selection state (a few MB on
 * the device) instead of scanning the buffers of all virtual ranks again, then enqueue their own next pass.  Same
 * stream as the leader's call, after it.
 */
int mc_stage_adopt(mc_handle* h, mc_handle* leader, const void* d_gathered, uint32_t n_ranks,
                   void* d_stage_out, int* done, void* stream);
/*
 * With mc_params.pass_group = g > 1 the adopting handles also ADD their narrowing-pass output to the leader's stage
 * buffer instead of filling their own (counts, histogram bins and candidate lists are additive): the GPU then hands
 * the other GPUs ONE buffer per pass -- 104 MB instead of g x 104 MB at 2521 rows.  mc_stage_shared(h) tells the host
 * which kind of buffer was just written: 1 = a shared pass output (gather the LEADER's buffer only; the next
 * mc_stage_next / mc_stage_adopt takes n_ranks / g), 0 = one buffer per virtual rank as before.
 */
int mc_stage_shared(mc_handle* h);

/*
 * Building blocks exposed for tests and for the callers either side of the path.
 */
/* z[t, i] (dtype) for paths path_offset.. : the shocks the simulator would draw */
int mc_debug_shocks(mc_handle* h, const mc_params* p, const void* d_pool, void* d_out, void* stream);
/* raw Philox-4x32-10 words: out[4*i..4*i+3] for counters ctr[4*i..], one key */
int mc_debug_philox(mc_handle* h, const uint32_t* d_ctr, uint32_t n, uint32_t key0, uint32_t key1,
                    uint32_t* d_out, void* stream);
/* np.percentile(x[r, :], pcts, method='linear') for r in rows of a [rows, ld] matrix */
int mc_percentiles(mc_handle* h, const void* d_x, int32_t dtype, uint64_t rows, uint64_t n,
                   uint64_t ld, const double* pcts, uint32_t n_pcts, double* h_out, void* stream);
/* Test knobs, per handle: capacity of the fp32 selection's side lists (< 0: automatic); no_side_lists bit 0: the lists
 * are not used, bit 1: mc_run keeps the terminal-value selection of a fan-chart request on the caller's stream instead
 * of running it on an internal stream beside the fan chart's passes;
 * process-wide: waves of blocks per launch of the non-persistent simulation kernels (>= 1, 0: default 24) and threads
 * per block of the fused GBM kernel (128 / 256 / 512, 0: default).  Invalid values select the default. */
int mc_debug_set_tuning(mc_handle* h, int side_cap, int no_side_lists, int waves, int nblock);
/* pilot size (values per row) of single-rank selections, read in place from the rows; 0: default (131072).  Process-wide. */
int mc_debug_set_pilot(uint32_t in_place_samples);
/* counters: 0 fused runs, 1 fused runs that had to be redone by the staged selection, 2 status of the last fused
 * run, 3 re-simulations of MC_FLAG_PATHS_VOLATILE runs */
int64_t mc_debug_counter(mc_handle* h, int which);
/* np.histogram(x, bins, range=(lo, hi)) on float64-cast values */
int mc_histogram(mc_handle* h, const void* d_x, int32_t dtype, uint64_t n, uint32_t bins,
                 double lo, double hi, uint64_t* h_counts, void* stream);

/*
 * np.histogram(x, bins, range=(lo, hi)) WITHOUT any allocation: mc_histogram for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  This entry point has no counterpart
 * in the reference.
 *
 * It needs no handle and calls no cudaMalloc / cudaFree / allocating library routine: the scratch the kernels need
 * (bin edges, running counts) lives in one file-scope __device__ array of 84 KB that every GPU loading the library
 * holds until the process exits (csrc/mc_static.cu).  Hence:
 *   - at most MC_STATIC_HIST_MAX_BINS bins; more is refused with MC_ERR_INVALID and a message in `err`;
 *   - every call on a device shares that array, which still holds the previous call's contents; the call's first
 *     kernel clears what must start at zero.  Calls that OVERLAP on the GPU (other streams, other threads, a graph
 *     replayed next to itself) race on it and give wrong counts: serialise them (one stream, or events);
 *   - the first call on a device asks the runtime for the array's device address (cudaGetSymbolAddress); make that
 *     call once before starting a capture.
 * Asynchronous: three small kernels on `stream`, no synchronisation, nothing copied to the host.
 *   d_x       [n] values (dtype MC_F32 / MC_F64), device memory of `device_ordinal`
 *   d_counts  [bins] uint64, device memory, overwritten
 *   err       caller's buffer for the error text (may be NULL), err_cap its size
 * Counts are those of mc_histogram on the same data, bit for bit.  The calling thread's current device is restored.
 */
#define MC_STATIC_HIST_MAX_BINS 4096u
int mc_histogram_static(int device_ordinal, const void* d_x, int32_t dtype, uint64_t n, uint32_t bins,
                        double lo, double hi, uint64_t* d_counts, void* stream, char* err, uint32_t err_cap);

/*
 * Input producers: what the routine reads, from raw close prices, for n_series price histories
 * in one launch.  Replaces, per history, the pandas producers of the two columns the routine reads --
 * `log_ret` (app.py:224) and `rvol_20` (app.py:233, sample std of the last 20 log returns x sqrt(252)) --, the
 * row dropping of app.py:392-393 (rows before `first_row` go: the reference's 200-day average, app.py:252, makes
 * that 199; rows before 20 always go, they have no rvol_20), and the routine's own first lines: `returns`
 * (app.py:592), `current_var` (app.py:606), `std_returns` (app.py:609); s0 is the call site's `current_price`
 * (app.py:1213).  All arithmetic is fp64; the pool is written in `pool_dtype` (what mc_run wants as d_pool).
 *   d_close   [n_series, close_ld] fp64, history b holds d_lengths[b] prices (all n_obs when d_lengths is NULL)
 *   d_pool    [n_series, pool_ld]: pool b = d_pool + b * pool_ld, d_inputs[b].pool_len entries
 *   d_inputs  [n_series] (device memory); pool_len < 100 is the reference's "return None" case (app.py:594-595)
 * Asynchronous on `stream`; nothing is copied to the host.
 */
#define MC_INPUT_NAN 1u          /* flags: a log return was NaN (non-positive or NaN price); the reference would drop the row */
typedef struct mc_series_inputs {
    double s0;                   /* close[last] */
    double var0;                 /* initial daily variance, mc_params.var0 */
    double pool_std;             /* np.std(returns), ddof = 0 */
    uint64_t pool_len;
    uint64_t flags;
} mc_series_inputs;
int mc_prepare_inputs(mc_handle* h, const double* d_close, const uint32_t* d_lengths, uint32_t n_series,
                      uint32_t n_obs, uint64_t close_ld, uint32_t first_row, int32_t pool_dtype, void* d_pool,
                      uint64_t pool_ld, mc_series_inputs* d_inputs, void* stream);

/*
 * The reference's model (bootstrap shocks + GARCH(1,1), app.py:612-657, fp64) for n_series price histories
 * in one pass: ONE simulation launch over (paths x histories), one exact selection over the [n_series, n_paths]
 * terminal matrix, one reduction launch, and (with MC_FLAG_FAN_CHART) one selection over the
 * n_series * (n_steps + 1) rows of the stored paths -- instead of ~20 launches per history.  Pools and
 * per-history scalars come from mc_prepare_inputs and are read from device memory.  History b draws from the
 * Philox key seed + b and steps exactly like mc_run in fp64, so its paths, percentiles and fan chart are
 * bit-identical to a single mc_run with that seed.  A history with fewer than `min_history` returns (100 in the
 * reference, app.py:594-595) or a NaN flag is skipped: h_results[b].n == 0.
 *   d_fit       [n_series] from mc_calibrate_garch: history b steps with its own omega/alpha/beta (NULL: the
 *               common ones of mc_batch_params); a history whose fit is not valid is skipped
 *   d_paths     [n_series, n_steps + 1, n_paths] fp64 (MC_FLAG_STORE_PATHS) or NULL
 *   d_terminal  [n_series, n_paths] fp64
 *   h_results   [n_series] (host); pct/pct_lo/pct_hi as in mc_run; sim_ms/total_ms/launches describe the batch
 *   h_fan       [n_series, n_steps + 1, n_path_pcts] fp64 (host) or NULL
 */
/*
 * GARCH(1,1) calibration (opt-in, it changes the results): per history, the (omega, alpha, beta)
 * that maximise the Gaussian likelihood of its returns under the reference's own recursion
 * var' = omega + alpha r^2 + beta var (app.py:632-633), in place of the constants of app.py:601-603.  Variance
 * targeting (omega = mean(r^2) (1 - alpha - beta)), deterministic grid refinement over (alpha, alpha + beta):
 * 16 x 16 candidates per level, `levels` levels (8 resolves both to ~1e-5).  Reads the pools and
 * mc_series_inputs that mc_prepare_inputs wrote (fp64 pool), writes d_fit[n_series] on the device; asynchronous.
 * `objective` = sum_t [log h_t + z_t^2 / h_t] on the standardised returns (-2 log-likelihood up to constants).
 */
typedef struct mc_garch_fit {
    double omega, alpha, beta;   /* in the units of the raw log returns: pass as mc_params.omega/alpha/beta */
    double objective;
    uint64_t valid;              /* 0: history skipped (too short / NaN) */
} mc_garch_fit;
int mc_calibrate_garch(mc_handle* h, const double* d_pool, uint64_t pool_ld, const mc_series_inputs* d_inputs,
                       uint32_t n_series, uint32_t min_history, uint32_t levels, mc_garch_fit* d_fit, void* stream);

typedef struct mc_batch_params {
    uint32_t struct_size;        /* sizeof(mc_batch_params) */
    uint32_t version;            /* MC_ABI_VERSION */
    uint64_t n_paths;            /* per history */
    uint64_t seed;
    uint32_t n_steps;
    uint32_t n_series;
    uint32_t min_history;
    uint32_t flags;              /* MC_FLAG_STORE_PATHS | MC_FLAG_FAN_CHART */
    double omega, alpha, beta;   /* app.py:601-603 */
    uint32_t n_pcts, n_path_pcts;
    double pcts[MC_MAX_PCTS];
    double path_pcts[MC_MAX_PATH_PCTS];
    double tail_pct;             /* < 0: none */
} mc_batch_params;
int mc_run_batch(mc_handle* h, const mc_batch_params* p, const void* d_pool, uint64_t pool_ld,
                 const mc_series_inputs* d_inputs, const mc_garch_fit* d_fit /* per-history omega/alpha/beta, or NULL */,
                 void* d_paths, void* d_terminal, mc_result* h_results, double* h_fan, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MC_B200_H */
