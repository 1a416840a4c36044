// mc_api.cu -- the C ABI of include/mc_b200.h: handle, workspace and the stage machine.
//
// The stage machine is what lets one code path serve both the single-GPU call
// (mc_run: every "gathered" buffer is the local stage buffer) and the path-sharded
// multi-GPU run (the host all-gathers each stage buffer over ranks between calls).
// Stages:   0        : moments header + pilot samples (terminal values, fan-chart rows)
//           1 .. P   : output of narrowing pass k (bracket counts, histograms, candidates)
//           P + 1    : tail (count, sum) + fixed-bin histogram counts
// P is normally 2; after the second pass the host reads a device counter of rows that
// still have undecided order statistics (identical on every rank) and keeps going if needed.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "mc_fused.cuh"
#include "mc_inputs.cuh"
#include "mc_select.cuh"
#include "mc_sim.cuh"
#include "mc_small.cuh"

using namespace mc;

namespace {
thread_local std::string g_create_error;

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

template <typename T>
struct PinnedBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMallocHost(&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};
struct SideBuf {
    DevBuf<float> values;
    DevBuf<unsigned int> counts;
    DevBuf<uint32_t> brackets;
    void release() { values.release(); counts.release(); brackets.release(); }
};

// Small results go to the host by stores into pinned (device-mapped) memory instead of copy-engine transfers: a
// cudaMemcpyAsync between two kernels costs the stream ~15 us (kernel -> copy engine -> kernel), a one-block
// kernel ~4 us (torch.profiler timeline of a config-2 job).
__global__ void words_to_host_kernel(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ a,
                                     size_t na, const unsigned long long* __restrict__ b, size_t nb,
                                     const unsigned long long* __restrict__ c, size_t nc) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < na + nb + nc; i += stride)
        dst[i] = i < na ? a[i] : (i < na + nb ? b[i - na] : c[i - na - nb]);
}
__global__ void u32_to_host_kernel(unsigned int* dst, const unsigned int* src, const unsigned int* src2) {
    *dst = *src + (src2 ? *src2 : 0u);
}
constexpr size_t HOST_STORE_MAX_BYTES = 512 * 1024;     // beyond this the copy engine wins
}  // namespace

struct mc_handle {
    int device = 0;
    DeviceInfo dev{};
    std::mutex mu;
    std::string err;

    DevBuf<BlockPartial> partials;
    DevBuf<float2> lut;                 // Box-Muller direction table of the fp32 normal shocks (mc_rng.cuh)
    DevBuf<DeviceResult> res;
    DevBuf<SelectPlan> plans;           // [2]: terminal, fan
    SelectPlan plan_shadow[2];          // what plans[slot] holds on the device ...
    bool plan_valid[2] = {false, false};
    cudaStream_t plan_stream[2] = {nullptr, nullptr};   // ... once this stream has reached the upload
    DevBuf<RowSel> states_t, states_f;
    DevBuf<ScanScratch> scratch_t, scratch_f;
    DevBuf<unsigned int> pending;       // [1] rows with undecided targets after the latest scan
    DevBuf<unsigned int> pending_t;     // [1] the same for the terminal problem while it runs on the side stream
    PinnedBuf<unsigned int> h_pending;
    DevBuf<double> pct_t;               // [3 * MC_MAX_PCTS]
    DevBuf<double> fan;                 // [rows * n_path_pcts]
    DevBuf<double> tail_scratch;        // [2 * 1024]
    DevBuf<double> edges;
    DevBuf<unsigned long long> hist_total;
    DevBuf<unsigned long long> stage;   // internal stage buffer (mc_run, mc_percentiles)
    SideBuf side_t, side_f;             // side lists of the fp32 selection passes (terminal / fan-chart rows)
    DevBuf<unsigned int> side_blocks;   // [1] blocks of the next pass that cannot use their side lists
    PinnedBuf<unsigned int> h_side_blocks;
    unsigned resimulations = 0;         // MC_FLAG_PATHS_VOLATILE: times the handle had to simulate again
    DevBuf<double> batch_pct;           // mc_run_batch: [3][n_series][n_pcts]
    DevBuf<mc_result> batch_res;        // mc_run_batch: [n_series]
    PinnedBuf<unsigned char> h_stage;   // pinned staging for the D2H of results
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev_pending = nullptr;   // recorded after the D2H copy of `pending`
    bool pending_unchecked = false;     // closing was enqueued before the host looked at `pending`

    // state of the run in flight
    bool active = false;
    mc_params p{};
    SelectProblem prob_t{}, prob_f{};
    bool has_t = false, has_f = false;
    int phase = 0;          // 0: gathered = stage 0; 1: gathered = a pass stage; 2: gathered = final stage
    int pass_idx = 0;       // narrowing passes enqueued so far
    int tail_idx = -1;
    size_t next_words = 0;  // words of the stage buffer most recently written
    unsigned launches = 0;
    void* d_terminal = nullptr;
    // what mc_stage_resimulate needs to run the simulation of the run in flight again
    SimBuffers sim_buffers{};
    bool has_pf = false;
    std::vector<double> pf_chol, pf_sigma, pf_drift, pf_s0, pf_w;

    // ---- fused closing (mc_fused.cuh): runs without a fan chart ----
    DevBuf<unsigned char> xchg;         // this handle's exchange buffer (layout `lay`)
    XchgLayout lay{};
    Peers peers{};                      // base[me] = xchg.p; the others are peers' buffers once attached
    bool peers_attached = false;
    std::vector<void*> ipc_opened;      // peer mappings to close
    DevBuf<FusedPlan> fplan;
    DevBuf<unsigned char> member;       // [FUSED_KB]
    DevBuf<ClosePartial> close_partials;
    DevBuf<unsigned long long> claim;   // [1]
    unsigned long long* last_pass_out = nullptr;   // where this handle's latest narrowing pass wrote (shared within a pass group)
    unsigned long long fused_seq = 0;   // sharded runs so far (the flags of a run carry its number)
    bool fused_pending = false;         // a fused run is enqueued; mc_fetch_result reads the pinned block
    bool fused_dirty = true;            // accumulators (own histogram slots, claim counter) need zeroing
    unsigned int edges_bins = 0;        // what h->edges currently holds
    double edges_lo = 0.0, edges_hi = 0.0;
    unsigned long long n_fused = 0, n_fused_fallback = 0;
    int last_fused_status = 0;
    // mc_run of a single-rank request with BOTH a fan chart and terminal percentiles: the terminal problem (one row,
    // ~0.2 ms of small kernels one behind the other) is enqueued on `side` and runs underneath the fan chart's passes
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool allow_overlap = false;         // set by mc_run around its stage calls (callers of the stage API own the stage buffers)
    bool overlap_t = false;             // this run: terminal kernels go to `side`
    // test knobs (mc_debug_set_tuning)
    int side_cap_override = -1;
    bool no_side_lists = false;
    bool no_overlap = false;
};

namespace {

int fail(mc_handle* h, int code, const std::string& msg) {
    if (h) h->err = msg; else g_create_error = msg;
    return code;
}
int cuda_fail(mc_handle* h, cudaError_t e, const char* where) {
    return fail(h, MC_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
// Every entry point works on its handle's device and leaves the calling thread's current device as it found it
// (a Streamlit session thread that asked for device 1 must not find itself on device 1 afterwards).
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess; else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};
#define MC_DEVICE(h)                                                      \
    DeviceGuard device_guard__((h)->device);                              \
    if (!device_guard__.ok) return fail((h), MC_ERR_CUDA, "cudaSetDevice failed")
#define MC_CUDA(h, call)                                                  \
    do {                                                                  \
        cudaError_t e__ = (call);                                         \
        if (e__ != cudaSuccess) return cuda_fail((h), e__, #call);        \
    } while (0)

// ---- stage buffer layouts (8-byte words) ----
struct Stage0Layout { size_t off_t, off_f, words; };
Stage0Layout layout_stage0(const mc_handle* h) {
    Stage0Layout L{};
    size_t w = STAGE0_HEADER;
    L.off_t = w;
    if (h->has_t) w += sample_words(h->prob_t);
    L.off_f = w;
    if (h->has_f) w += sample_words(h->prob_f);
    L.words = w;
    return L;
}
struct PassLayout { size_t off_t, off_f, words; };
PassLayout layout_pass(const mc_handle* h) {
    PassLayout L{};
    L.off_t = 0;
    size_t w = h->has_t ? pass_words(h->prob_t) : 0;
    L.off_f = w;
    if (h->has_f) w += pass_words(h->prob_f);
    L.words = w;
    return L;
}
size_t final_stage_words(const mc_params& p) { return 2 + (size_t)p.hist_bins; }

unsigned int ranks_of(const mc_params& p) { return p.n_ranks ? p.n_ranks : 1u; }
unsigned int pass_group_of(const mc_params& p) { return p.pass_group > 1 ? p.pass_group : 1u; }

// Side lists for an fp32 problem (mc_select.cuh: SideLists); a failed allocation just leaves them off.
void attach_side(mc_handle* h, SelectProblem& sp, SideBuf& buf) {
    sp.side = SideLists{};
    if (sp.dtype != MC_F32 || sp.n == 0 || sp.rows == 0 || h->no_side_lists) return;
    const unsigned long long chunks = select_chunks_f32(sp, h->dev.sm_count);
    unsigned int cap = side_list_cap(sp.n, chunks);
    if (h->side_cap_override >= 0) cap = (unsigned int)h->side_cap_override;                  // tests: force overflows
    if (cap == 0) return;
    if (buf.values.reserve(side_list_floats(sp.rows, chunks, cap)) != cudaSuccess ||
        buf.counts.reserve((size_t)sp.rows * chunks * SIDE_WARPS) != cudaSuccess ||
        buf.brackets.reserve((size_t)sp.rows * SIDE_BRACKET_WORDS) != cudaSuccess) {
        cudaGetLastError();
        buf.release();
        return;
    }
    sp.side.values = buf.values.p; sp.side.counts = buf.counts.p; sp.side.brackets = buf.brackets.p;
    sp.side.cap = cap; sp.side.chunks = (unsigned int)chunks;
}

void setup_sampling(SelectProblem& sp, unsigned int n_ranks, bool contiguous) {
    sp.sample_contiguous = contiguous ? 1 : 0;
    sp.pilot_in_place = n_ranks <= 1 ? 1 : 0;
    if (sp.pilot_in_place) {            // one rank: the brackets are taken from the rows themselves
        // one row (terminal values): a single block reads the pilot, and pass 1 over one row is cheap anyway
        const unsigned int m = sp.rows == 1 ? (unsigned int)SEL_SAMPLE_MAX : pilot_in_place_size(contiguous);
        sp.m_slot = 0;
        sp.m_local = sp.n < m ? (unsigned int)sp.n : m;
        return;
    }
    sp.m_slot = sample_slot(sp.dtype, n_ranks);
    sp.m_local = sp.n < sp.m_slot ? (unsigned int)sp.n : sp.m_slot;
}

int validate(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool, const void* d_shocks,
             void* d_paths, void* d_terminal) {
    if (!p) return fail(h, MC_ERR_INVALID, "params is NULL");
    if (p->struct_size != sizeof(mc_params) || p->version != MC_ABI_VERSION)
        return fail(h, MC_ERR_VERSION, "mc_params struct_size/version mismatch");
    if (p->n_paths_global == 0 || p->n_paths > p->n_paths_global)
        return fail(h, MC_ERR_INVALID, "n_paths_global must be >= n_paths and > 0");
    if (p->n_paths >= (1ull << 32)) return fail(h, MC_ERR_INVALID, "n_paths per call must be < 2^32");
    if (p->n_steps == 0 || p->n_steps > 1000000u) return fail(h, MC_ERR_INVALID, "n_steps out of range");
    if (p->dtype != MC_F32 && p->dtype != MC_F64) return fail(h, MC_ERR_INVALID, "bad dtype");
    if (p->shock < 0 || p->shock > MC_SHOCK_INJECTED) return fail(h, MC_ERR_INVALID, "bad shock mode");
    if (p->var_model != MC_VAR_GARCH && p->var_model != MC_VAR_CONST) return fail(h, MC_ERR_INVALID, "bad var model");
    if (p->n_assets < 1 || p->n_assets > MC_MAX_ASSETS) return fail(h, MC_ERR_INVALID, "n_assets out of range");
    if (!(p->flags & MC_FLAG_TERMINAL_GIVEN) && p->n_assets > 1 && (!pf || !pf->h_chol || !pf->h_sigma_step || !pf->h_drift_step || !pf->h_s0 || !pf->h_weights))
        return fail(h, MC_ERR_INVALID, "portfolio description missing");
    if (p->n_assets > 1 && (p->shock == MC_SHOCK_BOOTSTRAP || p->var_model != MC_VAR_CONST))
        return fail(h, MC_ERR_INVALID, "portfolio mode supports normal/injected shocks with constant variance");
    const bool simulates = !(p->flags & MC_FLAG_TERMINAL_GIVEN);
    if (simulates && p->shock == MC_SHOCK_BOOTSTRAP && (!d_pool || p->pool_len == 0))
        return fail(h, MC_ERR_INVALID, "bootstrap needs a pool");
    if (simulates && p->shock == MC_SHOCK_INJECTED && !d_shocks) return fail(h, MC_ERR_INVALID, "injected shocks missing");
    if (!d_terminal && p->n_paths) return fail(h, MC_ERR_INVALID, "d_terminal is NULL");
    if ((p->flags & MC_FLAG_STORE_PATHS) && !d_paths && p->n_paths) return fail(h, MC_ERR_INVALID, "d_paths is NULL");
    if ((p->flags & MC_FLAG_FAN_CHART) && !(p->flags & MC_FLAG_STORE_PATHS))
        return fail(h, MC_ERR_INVALID, "fan chart needs stored paths");
    if (p->paths_ld && p->paths_ld < p->n_paths) return fail(h, MC_ERR_INVALID, "paths_ld < n_paths");
    if (p->n_pcts > MC_MAX_PCTS || p->n_path_pcts > MC_MAX_PATH_PCTS) return fail(h, MC_ERR_INVALID, "too many percentiles");
    if (p->hist_bins > (1u << 16)) return fail(h, MC_ERR_INVALID, "hist_bins > 65536");
    if (!(p->var0 >= 0.0) || !(p->s0 == p->s0)) return fail(h, MC_ERR_INVALID, "var0 must be >= 0");
    if (p->pass_group > 1 && (ranks_of(*p) % p->pass_group) != 0)
        return fail(h, MC_ERR_INVALID, "pass_group must divide n_ranks");
    if ((p->flags & MC_FLAG_FULL_RES_NORMALS) && p->n_assets > 1 && p->dtype == MC_F32 && p->shock == MC_SHOCK_NORMAL)
        return fail(h, MC_ERR_INVALID, "MC_FLAG_FULL_RES_NORMALS: single-asset kernels only");
    if ((p->flags & MC_FLAG_TERMINAL_GIVEN) && (p->flags & (MC_FLAG_STORE_PATHS | MC_FLAG_FAN_CHART)))
        return fail(h, MC_ERR_INVALID, "MC_FLAG_TERMINAL_GIVEN reduces terminal values only");
    return MC_OK;
}

int upload_plan(mc_handle* h, SelectProblem& sp, int slot, cudaStream_t s) {
    sp.d_plan = h->plans.p + slot;
    // repeated calls with the same targets on the same stream (the dashboard's case) leave the device copy as it is
    if (h->plan_valid[slot] && h->plan_stream[slot] == s && memcmp(&h->plan_shadow[slot], &sp.plan, sizeof(SelectPlan)) == 0) return MC_OK;
    // pageable host -> device async copy is staged by the runtime before returning
    MC_CUDA(h, cudaMemcpyAsync(sp.d_plan, &sp.plan, sizeof(SelectPlan), cudaMemcpyHostToDevice, s));
    memcpy(&h->plan_shadow[slot], &sp.plan, sizeof(SelectPlan));
    h->plan_valid[slot] = true;
    h->plan_stream[slot] = s;
    return MC_OK;
}

// np.percentile(x[r, :], pcts) for `rows` rows of a device matrix, results left on the device:
// d_out[rows][n_pcts] (and the two order statistics each one interpolates, if wanted).  Synchronises the
// stream once per narrowing pass after the first (the host reads the count of undecided rows).
int select_rows(mc_handle* h, const void* d_x, int dtype, uint64_t rows, uint64_t n, uint64_t ld, const double* pcts,
                uint32_t n_pcts, bool iid_columns, double* d_out, double* d_lo, double* d_hi, unsigned* launches,
                cudaStream_t s) {
    SelectProblem sp{};
    if (!build_select_plan(pcts, (int)n_pcts, n, dtype, &sp.plan))
        return fail(h, MC_ERR_INVALID, "percentiles must be ascending and within [0, 100]");
    sp.x = d_x; sp.dtype = dtype; sp.rows = rows; sp.n = n; sp.ld = ld;
    setup_sampling(sp, 1, iid_columns);                // arbitrary data: strided sample
    attach_side(h, sp, h->side_f);
    MC_CUDA(h, h->states_f.reserve(rows));
    MC_CUDA(h, h->scratch_f.reserve(rows));
    sp.d_states = h->states_f.p;
    sp.d_scratch = h->scratch_f.p;
    int rc = upload_plan(h, sp, 1, s);
    if (rc) return rc;
    const size_t w0 = STAGE0_HEADER + sample_words(sp), wk = pass_words(sp);
    MC_CUDA(h, h->stage.reserve(w0 > wk ? w0 : wk));
    unsigned long long* st = h->stage.p;
    unsigned count = 0;
    MC_CUDA(h, launch_reduce_partials(h->partials.p, 0, st, 0, sp.m_local, s));
    MC_CUDA(h, launch_select_sample(sp, st + STAGE0_HEADER, s));
    MC_CUDA(h, launch_select_brackets(sp, st, STAGE0_HEADER, 7, 1, w0, h->dev.max_smem_optin, s));
    count += sp.pilot_in_place ? 2 : 3;       // kernels only (the pilot is read in place: no sampling kernel)
    for (int pass = 0;; ++pass) {
        MC_CUDA(h, launch_select_pass(sp, pass == 0, st, h->dev.sm_count, s));
        MC_CUDA(h, cudaMemsetAsync(h->pending.p, 0, sizeof(unsigned int), s));
        MC_CUDA(h, launch_select_scan(sp, st, 1, wk, h->pending.p, s));
        count += 3;
        if (pass == 0) continue;
        MC_CUDA(h, cudaMemcpyAsync(h->h_pending.p, h->pending.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, s));
        MC_CUDA(h, cudaStreamSynchronize(s));
        if (*h->h_pending.p == 0) break;
        if (pass + 1 >= SEL_MAX_PASSES) return fail(h, MC_ERR_STATE, "selection did not converge");
    }
    MC_CUDA(h, launch_select_finish(sp, d_out, d_lo, d_hi, s));
    if (launches) *launches += count + 1;
    return MC_OK;
}

// ------------------------------------------------------------------ fused closing: host side
constexpr unsigned int PEER_CAND_BYTES = 16u << 20;       // a rank's candidate lists, fixed once peers are attached

bool fused_eligible(const mc_params& p) {
    if (p.flags & (MC_FLAG_FAN_CHART | MC_FLAG_PATHS_VOLATILE | MC_FLAG_TERMINAL_GIVEN)) return false;
    if (p.hist_bins && (!(p.hist_lo < p.hist_hi) || p.hist_bins > (unsigned)FUSED_MAX_NPH_BINS)) return false;   // range from the data: general path
    if (ranks_of(p) > (unsigned)FUSED_MAX_RANKS) return false;
    return true;
}

// The window of terminal values the key histogram resolves: s0 e^(drift T) times e^(+-W spread), where spread is
// the standard deviation of the log return the model itself implies.  Only its placement depends on this
// estimate, never a result: values outside it land in the two open-ended end bins, and a target that falls
// into an over-full bin sends the run to the general selection.
KeyWindow key_window(const mc_params& p, const mc_portfolio* pf) {
    const double T = (double)p.n_steps;
    double lo_ln, hi_ln;
    if (p.n_assets > 1 && pf) {
        double smax = 0.0, dmin = INFINITY, dmax = -INFINITY;
        for (unsigned i = 0; i < p.n_assets; ++i) {
            smax = fmax(smax, fabs(pf->h_sigma_step[i]));
            dmin = fmin(dmin, pf->h_drift_step[i]);
            dmax = fmax(dmax, pf->h_drift_step[i] + 0.5 * pf->h_sigma_step[i] * pf->h_sigma_step[i]);
        }
        const double spread = smax * sqrt(T), c = log(fabs(p.s0) > 0 ? fabs(p.s0) : 1.0);
        lo_ln = c + T * dmin - 6.5 * spread;
        hi_ln = c + T * dmax + 6.5 * spread;
    } else {
        double total = 0.0;
        if (p.var_model == MC_VAR_CONST) {
            total = p.var0 * T;
        } else {                                   // E v_{t+1} = omega + (alpha + beta) E v_t  (E z^2 ~ 1: app.py:609)
            double v = p.var0;
            const double phi = p.alpha + p.beta;
            for (unsigned t = 0; t < p.n_steps; ++t) { total += v; v = p.omega + phi * v; if (!(v < 1e6)) v = 1e6; }
        }
        const double spread = sqrt(total > 0 ? total : 0.0), W = p.var_model == MC_VAR_CONST ? 6.5 : 8.0;
        const double c = log(fabs(p.s0) > 0 ? fabs(p.s0) : 1.0) + T * p.drift;
        lo_ln = c - W * spread;
        hi_ln = c + W * spread;
    }
    if (!(lo_ln == lo_ln) || !(hi_ln == hi_ln)) { lo_ln = -80.0; hi_ln = 80.0; }
    lo_ln = fmax(lo_ln, -80.0); hi_ln = fmin(hi_ln, 80.0);
    if (hi_ln < lo_ln) hi_ln = lo_ln;
    const double lo = exp(lo_ln), hi = exp(hi_ln);
    return p.dtype == MC_F64 ? make_window(key32_of(lo), key32_of(hi)) : make_window(key32_of((float)lo), key32_of((float)hi));
}

// (re)allocate this handle's exchange buffer; once peers are attached it must not move
int ensure_xchg(mc_handle* h, unsigned int n_ranks, unsigned int cand_bytes) {
    if (h->xchg.p && h->lay.n_ranks >= n_ranks && h->lay.cand_bytes >= cand_bytes) return MC_OK;
    if (h->peers_attached) return fail(h, MC_ERR_STATE, "exchange buffer is attached to peers and cannot grow");
    const XchgLayout L = xchg_layout(n_ranks, cand_bytes);
    MC_CUDA(h, h->xchg.reserve(L.total));
    MC_CUDA(h, cudaMemset(h->xchg.p, 0, L.total));
    h->lay = L;
    h->peers = Peers{};
    h->peers.base[0] = h->xchg.p;
    h->peers.n_ranks = 1; h->peers.me = 0; h->peers.seq = 0;
    h->fused_dirty = true;
    return MC_OK;
}

// workspace of the fused closing for up to `bins` histogram bins (allocations may wait for the device: a sharded
// run must find everything in place, so mc_peer_export reserves for the largest request)
int fused_reserve(mc_handle* h, unsigned int bins) {
    MC_CUDA(h, h->fplan.reserve(1));
    MC_CUDA(h, h->member.reserve(FUSED_KB));
    MC_CUDA(h, h->close_partials.reserve(8 * (size_t)h->dev.sm_count + 8));
    MC_CUDA(h, h->claim.reserve(1));
    MC_CUDA(h, h->h_stage.reserve(sizeof(DeviceResult) + (size_t)bins * 8));
    if (bins) {
        const double* before = h->edges.p;
        MC_CUDA(h, h->edges.reserve(2 * ((size_t)bins + 2)));
        if (h->edges.p != before) h->edges_bins = 0;
        MC_CUDA(h, h->hist_total.reserve(bins));
    }
    return MC_OK;
}

// enqueue a whole run with the fused closing; the result is read later by fused_fetch
int fused_enqueue(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool, const void* d_shocks,
                  void* d_paths, void* d_terminal, bool sharded, cudaStream_t s) {
    int rc = validate(h, p, pf, d_pool, d_shocks, d_paths, d_terminal);
    if (rc) return rc;
    MC_DEVICE(h);
    const unsigned int R = ranks_of(*p);
    if (sharded) {
        if (!h->peers_attached || h->peers.n_ranks != R) return fail(h, MC_ERR_STATE, "mc_run_sharded: peers not attached for this many ranks");
    } else {
        if (R != 1) return fail(h, MC_ERR_INVALID, "single-rank call with n_ranks > 1");
        if (h->peers_attached && h->peers.n_ranks != 1) return fail(h, MC_ERR_STATE, "handle is attached to peers: use mc_run_sharded");
        unsigned long long want = p->n_paths * (p->dtype == MC_F64 ? 8ull : 4ull) / 8ull;
        if (want < (1ull << 20)) want = 1ull << 20;
        if (want > (1ull << 30)) want = 1ull << 30;
        want = (want + 255ull) & ~255ull;
        if ((rc = ensure_xchg(h, 1, (unsigned int)want))) return rc;
    }
    h->active = false;
    h->p = *p;
    h->launches = 0;
    h->has_t = p->n_pcts > 0 && !(p->flags & MC_FLAG_NO_TERMINAL_SELECT);
    h->has_f = false;
    h->tail_idx = -1;
    h->d_terminal = d_terminal;
    h->prob_t = SelectProblem{};
    if (h->has_t) {
        if (!build_select_plan(p->pcts, (int)p->n_pcts, p->n_paths_global, p->dtype, &h->prob_t.plan))
            return fail(h, MC_ERR_INVALID, "percentiles must be ascending and within [0, 100]");
        if (p->tail_pct >= 0.0) {
            for (unsigned k = 0; k < p->n_pcts; ++k) if (p->pcts[k] == p->tail_pct) h->tail_idx = (int)k;
            if (h->tail_idx < 0) return fail(h, MC_ERR_INVALID, "tail_pct must be one of pcts");
        }
    } else {
        build_select_plan(nullptr, 0, p->n_paths_global, p->dtype, &h->prob_t.plan);
    }
    if ((rc = upload_plan(h, h->prob_t, 0, s))) return rc;
    if ((rc = fused_reserve(h, p->hist_bins))) return rc;
    if (p->hist_bins) {
        if (h->edges_bins != p->hist_bins || h->edges_lo != p->hist_lo || h->edges_hi != p->hist_hi) {
            MC_CUDA(h, launch_hist_edges(p->hist_bins, p->hist_lo, p->hist_hi, h->res.p, h->edges.p, s));
            h->edges_bins = p->hist_bins; h->edges_lo = p->hist_lo; h->edges_hi = p->hist_hi;
            h->launches += 1;
        }
    }
    unsigned char* mine = h->peers.base[h->peers.me];
    unsigned int* my_hist = reinterpret_cast<unsigned int*>(mine + h->lay.off_hist) + (size_t)h->peers.me * FUSED_KB;
    unsigned int* my_nph = reinterpret_cast<unsigned int*>(mine + h->lay.off_nph) + (size_t)h->peers.me * FUSED_MAX_NPH_BINS;
    if (h->fused_dirty) {          // normally the last block of the previous run left these at zero
        MC_CUDA(h, cudaMemsetAsync(my_hist, 0, (size_t)FUSED_KB * 4, s));
        MC_CUDA(h, cudaMemsetAsync(my_nph, 0, (size_t)FUSED_MAX_NPH_BINS * 4, s));
        MC_CUDA(h, cudaMemsetAsync(h->claim.p, 0, sizeof(unsigned long long), s));
    }
    h->fused_dirty = true;         // until this run has been enqueued completely

    FusedArgs fa{};
    fa.x = d_terminal; fa.dtype = p->dtype; fa.n = p->n_paths; fa.center = p->s0;
    fa.win = key_window(*p, pf);
    fa.lay = h->lay;
    fa.peers = h->peers;
    if (sharded) fa.peers.seq = ++h->fused_seq;
    fa.d_plan = h->prob_t.d_plan;
    fa.n_targets = h->prob_t.plan.n_targets; fa.n_pcts = h->has_t ? (int)p->n_pcts : 0; fa.tail_idx = h->tail_idx;
    fa.plan = h->fplan.p; fa.member = h->member.p; fa.partials = h->close_partials.p;
    fa.nph_bins = p->hist_bins; fa.edges = h->edges.p; fa.hist_lo = p->hist_lo; fa.hist_hi = p->hist_hi;
    fa.res = h->res.p; fa.hist_total = h->hist_total.p; fa.host_out = h->h_stage.p; fa.claim = h->claim.p;

    SinkArgs sink{};
    const bool in_sim = sim_takes_sink(*p) && p->n_paths > 0;
    if (in_sim) {
        sink.win = fa.win; sink.g_khist = my_hist; sink.claim = h->claim.p;
        if (p->hist_bins && p->hist_bins <= (unsigned)FUSED_SINK_NPH_BINS) {
            sink.g_nphist = my_nph; sink.nph_bins = p->hist_bins;
            sink.nph_edges = reinterpret_cast<const float*>(h->edges.p + p->hist_bins + 1);
            fa.nph_in_sim = 1;
        }
    }
    MC_CUDA(h, cudaEventRecord(h->ev[0], s));
    SimBuffers b{};
    b.d_pool = d_pool; b.d_shocks = d_shocks; b.d_paths = d_paths; b.d_terminal = d_terminal;
    b.d_partials = h->partials.p; b.d_lut = h->lut.p; b.sink = in_sim ? &sink : nullptr;
    int grid = 0;
    if (p->n_paths) {
        MC_CUDA(h, launch_simulation(*p, pf, b, h->dev, &grid, s));
        h->launches += 1;
    }
    MC_CUDA(h, cudaEventRecord(h->ev[1], s));
    if (!in_sim && p->n_paths) {
        MC_CUDA(h, launch_terminal_hist(fa, h->dev.sm_count, s));
        h->launches += 1;
    }
    MC_CUDA(h, launch_fused_plan(fa, s));
    MC_CUDA(h, launch_fused_closing(fa, h->dev.sm_count, &grid, s));
    MC_CUDA(h, launch_fused_select(fa, s));
    h->launches += 3;
    MC_CUDA(h, cudaEventRecord(h->ev[2], s));
    h->fused_dirty = false;
    h->fused_pending = true;
    h->n_fused += 1;
    return MC_OK;
}

// wait for the fused run and unpack the pinned result block; *status = FUSED_* of the run
int fused_fetch(mc_handle* h, mc_result* r, uint64_t* h_hist, cudaStream_t s, int* status) {
    MC_DEVICE(h);
    MC_CUDA(h, cudaStreamSynchronize(s));
    h->fused_pending = false;
    const mc_params& p = h->p;
    DeviceResult d;
    memcpy(&d, h->h_stage.p, sizeof(d));
    *status = (int)d.status;
    h->last_fused_status = *status;
    memset(r, 0, sizeof(*r));
    r->n = d.n; r->n_above = d.n_above; r->mean = d.mean; r->std = d.std; r->min = d.vmin; r->max = d.vmax;
    if (h->has_t) {
        memcpy(r->pct, d.pct, sizeof(d.pct));
        memcpy(r->pct_lo, d.pct_lo, sizeof(d.pct_lo));
        memcpy(r->pct_hi, d.pct_hi, sizeof(d.pct_hi));
    }
    r->tail_count = d.tail_count; r->tail_sum = d.tail_sum;
    r->hist_lo = d.hist_lo; r->hist_hi = d.hist_hi;
    if (p.hist_bins && h_hist) memcpy(h_hist, h->h_stage.p + sizeof(DeviceResult), (size_t)p.hist_bins * 8);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]) == cudaSuccess) r->sim_ms = ms;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[2]) == cudaSuccess) r->total_ms = ms;
    r->launches = h->launches;
    r->status = (uint32_t)*status;
    if (*status != FUSED_OK) { h->n_fused_fallback += 1; h->fused_dirty = true; }
    if (*status == FUSED_PEER_TIMEOUT) return fail(h, MC_ERR_STATE, "a peer rank never delivered its part of the exchange (20 s)");
    return MC_OK;
}

// ------------------------------------------------------------------ the reference's own call sizes: two launches
bool small_eligible(const mc_params& p) {
    if (!(p.flags & MC_FLAG_FAN_CHART) || !(p.flags & MC_FLAG_STORE_PATHS) || p.n_path_pcts == 0) return false;
    if (p.flags & (MC_FLAG_PATHS_VOLATILE | MC_FLAG_TERMINAL_GIVEN)) return false;
    if (p.hist_bins || ranks_of(p) > 1 || p.n_paths != p.n_paths_global) return false;
    if (p.n_paths == 0 || p.n_paths > (p.dtype == MC_F64 ? SMALL_MAX_PATHS_F64 : SMALL_MAX_PATHS_F32)) return false;
    return (size_t)p.n_steps + 1 <= 65535;
}

// returns MC_OK with *done = false when the job turned out not to fit this path after all (nothing is lost: the
// caller runs the general path)
int small_run(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool, const void* d_shocks,
              void* d_paths, void* d_terminal, mc_result* r, double* h_fan, cudaStream_t s, bool* done) {
    *done = false;
    int rc = validate(h, p, pf, d_pool, d_shocks, d_paths, d_terminal);
    if (rc) return rc;
    MC_DEVICE(h);
    h->active = false;
    h->fused_pending = false;
    h->p = *p;
    h->launches = 0;
    h->has_t = p->n_pcts > 0 && !(p->flags & MC_FLAG_NO_TERMINAL_SELECT);
    h->has_f = true;
    h->tail_idx = -1;
    h->d_terminal = d_terminal;
    h->prob_t = SelectProblem{};
    h->prob_f = SelectProblem{};
    if (h->has_t) {
        if (!build_select_plan(p->pcts, (int)p->n_pcts, p->n_paths_global, p->dtype, &h->prob_t.plan))
            return fail(h, MC_ERR_INVALID, "percentiles must be ascending and within [0, 100]");
        if (p->tail_pct >= 0.0) {
            for (unsigned k = 0; k < p->n_pcts; ++k) if (p->pcts[k] == p->tail_pct) h->tail_idx = (int)k;
            if (h->tail_idx < 0) return fail(h, MC_ERR_INVALID, "tail_pct must be one of pcts");
        }
    } else {
        build_select_plan(nullptr, 0, p->n_paths_global, p->dtype, &h->prob_t.plan);
    }
    if (!build_select_plan(p->path_pcts, (int)p->n_path_pcts, p->n_paths_global, p->dtype, &h->prob_f.plan))
        return fail(h, MC_ERR_INVALID, "path percentiles must be ascending and within [0, 100]");
    if ((rc = upload_plan(h, h->prob_t, 0, s))) return rc;
    if ((rc = upload_plan(h, h->prob_f, 1, s))) return rc;
    const size_t rows = (size_t)p->n_steps + 1, fan_n = rows * p->n_path_pcts;
    MC_CUDA(h, h->fan.reserve(fan_n));
    MC_CUDA(h, h->h_stage.reserve(sizeof(DeviceResult) + fan_n * 8));
    MC_CUDA(h, cudaEventRecord(h->ev[0], s));
    SimBuffers b{};
    b.d_pool = d_pool; b.d_shocks = d_shocks; b.d_paths = d_paths; b.d_terminal = d_terminal;
    b.d_partials = h->partials.p; b.d_lut = h->lut.p; b.sink = nullptr;
    int grid = 0;
    MC_CUDA(h, launch_simulation(*p, pf, b, h->dev, &grid, s));
    MC_CUDA(h, cudaEventRecord(h->ev[1], s));
    if (grid > SMALL_MAX_PARTIALS) return MC_OK;          // cannot happen at these sizes; the general path redoes the job
    SmallArgs a{};
    a.paths = d_paths; a.dtype = p->dtype; a.n = p->n_paths; a.ld = p->paths_ld ? p->paths_ld : p->n_paths;
    a.rows = (unsigned int)rows;
    a.plan_t = h->prob_t.d_plan; a.plan_f = h->prob_f.d_plan;
    a.n_pcts = h->has_t ? (int)p->n_pcts : 0; a.n_path_pcts = (int)p->n_path_pcts; a.tail_idx = h->tail_idx;
    a.partials = h->partials.p; a.n_partials = grid; a.center = p->s0;
    a.res = h->res.p; a.fan = h->fan.p; a.host_out = h->h_stage.p;
    a.bad = h->h_pending.p;
    *h->h_pending.p = 0u;
    MC_CUDA(h, launch_small_rows(a, s));
    MC_CUDA(h, cudaEventRecord(h->ev[2], s));
    MC_CUDA(h, cudaStreamSynchronize(s));
    if (*h->h_pending.p != 0u) return MC_OK;               // never seen; the general path redoes the job
    DeviceResult d;
    memcpy(&d, h->h_stage.p, sizeof(d));
    memset(r, 0, sizeof(*r));
    r->n = d.n; r->n_above = d.n_above; r->mean = d.mean; r->std = d.std; r->min = d.vmin; r->max = d.vmax;
    if (h->has_t) {
        memcpy(r->pct, d.pct, sizeof(d.pct));
        memcpy(r->pct_lo, d.pct_lo, sizeof(d.pct_lo));
        memcpy(r->pct_hi, d.pct_hi, sizeof(d.pct_hi));
    }
    r->tail_count = d.tail_count; r->tail_sum = d.tail_sum;
    if (h_fan) memcpy(h_fan, h->h_stage.p + sizeof(DeviceResult), fan_n * 8);       // already [pct][row]
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]) == cudaSuccess) r->sim_ms = ms;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[2]) == cudaSuccess) r->total_ms = ms;
    r->launches = 2;
    *done = true;
    return MC_OK;
}

}  // namespace

extern "C" {

int mc_abi_version(void) { return (int)MC_ABI_VERSION; }

const char* mc_last_error(const mc_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int mc_create(int device_ordinal, mc_handle** out) {
    if (!out) return fail(nullptr, MC_ERR_INVALID, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceCount");
    if (device_ordinal < 0 || device_ordinal >= count) return fail(nullptr, MC_ERR_INVALID, "no such CUDA device");
    mc_handle* h = new (std::nothrow) mc_handle();
    if (!h) return fail(nullptr, MC_ERR_INVALID, "out of host memory");
    h->device = device_ordinal;
    cudaDeviceProp prop{};
    DeviceGuard guard(device_ordinal);
    if ((e = guard.ok ? cudaSuccess : cudaErrorInvalidDevice) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, device_ordinal)) != cudaSuccess) {
        delete h;
        return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
    }
    if (prop.major < 10) {
        delete h;
        return fail(nullptr, MC_ERR_INVALID, "this library is built for sm_100a (Blackwell) only");
    }
    h->dev.sm_count = prop.multiProcessorCount;
    h->dev.max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    bool ok = h->partials.reserve(MAX_SIM_GRID) == cudaSuccess && h->res.reserve(1) == cudaSuccess &&
              h->plans.reserve(2) == cudaSuccess && h->states_t.reserve(1) == cudaSuccess &&
              h->scratch_t.reserve(1) == cudaSuccess &&
              h->pct_t.reserve(3 * MC_MAX_PCTS) == cudaSuccess && h->tail_scratch.reserve(2 * 1024) == cudaSuccess &&
              h->pending.reserve(1) == cudaSuccess && h->pending_t.reserve(1) == cudaSuccess && h->h_pending.reserve(1) == cudaSuccess &&
              h->side_blocks.reserve(1) == cudaSuccess && h->h_side_blocks.reserve(1) == cudaSuccess &&
              h->lut.reserve(4096) == cudaSuccess && launch_build_direction_table(h->lut.p, nullptr) == cudaSuccess &&
              cudaStreamSynchronize(nullptr) == cudaSuccess;
    for (int i = 0; i < 3 && ok; ++i) ok = cudaEventCreate(&h->ev[i]) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&h->ev_pending, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) == cudaSuccess &&
         cudaMemset(h->pending_t.p, 0, sizeof(unsigned int)) == cudaSuccess;
    if (!ok) {
        e = cudaGetLastError();
        mc_destroy(h);
        return cuda_fail(nullptr, e, "workspace allocation");
    }
    *out = h;
    return MC_OK;
}

int mc_destroy(mc_handle* h) {
    if (!h) return MC_OK;
    DeviceGuard guard(h->device);
    h->partials.release(); h->lut.release(); h->res.release(); h->plans.release(); h->states_t.release(); h->states_f.release(); h->scratch_t.release(); h->scratch_f.release();
    h->pct_t.release(); h->fan.release(); h->tail_scratch.release(); h->edges.release(); h->hist_total.release();
    h->batch_pct.release(); h->batch_res.release(); h->side_t.release(); h->side_f.release();
    h->side_blocks.release(); h->h_side_blocks.release();
    for (void* q : h->ipc_opened) cudaIpcCloseMemHandle(q);
    h->xchg.release(); h->fplan.release(); h->member.release(); h->close_partials.release(); h->claim.release();
    h->stage.release(); h->h_stage.release(); h->pending.release(); h->pending_t.release(); h->h_pending.release();
    if (h->side) { cudaStreamSynchronize(h->side); cudaStreamDestroy(h->side); }
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    for (auto& ev : h->ev) if (ev) cudaEventDestroy(ev);
    if (h->ev_pending) cudaEventDestroy(h->ev_pending);
    delete h;
    return MC_OK;
}

int64_t mc_stage_max_words(const mc_params* p) {
    if (!p) return -1;
    const unsigned int R = ranks_of(*p);
    const size_t ku = (size_t)key_units(p->dtype), slot = R <= 1 ? 0 : sample_slot(p->dtype, R);      // one rank: pilot read in place
    const size_t rows_f = ((p->flags & MC_FLAG_FAN_CHART) && p->n_path_pcts) ? (size_t)p->n_steps + 1 : 0;
    const bool sel_t = p->n_pcts > 0 && !(p->flags & MC_FLAG_NO_TERMINAL_SELECT);
    const size_t nt_t = sel_t ? 2 * (size_t)p->n_pcts : 0, nt_f = 2 * (size_t)p->n_path_pcts;
    size_t w0 = STAGE0_HEADER + (sel_t ? (slot * ku + 1) / 2 : 0) + (rows_f * slot * ku + 1) / 2;
    size_t wk = (row_slot_u32(p->dtype, (int)nt_t) + rows_f * row_slot_u32(p->dtype, (int)nt_f)) / 2;
    size_t wf = 2 + (size_t)p->hist_bins;
    size_t m = w0 > wk ? w0 : wk;
    return (int64_t)(m > wf ? m : wf);
}

int64_t mc_stage_words(const mc_handle* h) { return h ? (int64_t)h->next_words : -1; }

static int stage_begin_locked(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool,
                              const void* d_shocks, void* d_paths, void* d_terminal, void* d_stage_out,
                              uint64_t cap_words, cudaStream_t s) {
    int rc = validate(h, p, pf, d_pool, d_shocks, d_paths, d_terminal);
    if (rc) return rc;
    MC_DEVICE(h);
    h->active = false;
    h->fused_pending = false;
    h->p = *p;
    h->last_pass_out = nullptr;
    h->launches = 0;
    h->phase = 0;
    h->pass_idx = 0;
    h->pending_unchecked = false;
    h->d_terminal = d_terminal;
    h->has_t = p->n_pcts > 0 && !(p->flags & MC_FLAG_NO_TERMINAL_SELECT);
    h->has_f = (p->flags & MC_FLAG_FAN_CHART) && p->n_path_pcts > 0;
    h->tail_idx = -1;
    const unsigned int R = ranks_of(*p);
    if (h->has_t) {
        if (!build_select_plan(p->pcts, (int)p->n_pcts, p->n_paths_global, p->dtype, &h->prob_t.plan))
            return fail(h, MC_ERR_INVALID, "percentiles must be ascending and within [0, 100]");
        h->prob_t.x = d_terminal; h->prob_t.dtype = p->dtype; h->prob_t.rows = 1; h->prob_t.n = p->n_paths;
        h->prob_t.ld = p->n_paths; h->prob_t.d_states = h->states_t.p; h->prob_t.d_scratch = h->scratch_t.p;
        setup_sampling(h->prob_t, R, true);            // paths are iid: the first values are a random sample
        attach_side(h, h->prob_t, h->side_t);
        if ((rc = upload_plan(h, h->prob_t, 0, s))) return rc;
        if (p->tail_pct >= 0.0) {
            for (unsigned k = 0; k < p->n_pcts; ++k) if (p->pcts[k] == p->tail_pct) h->tail_idx = (int)k;
            if (h->tail_idx < 0) return fail(h, MC_ERR_INVALID, "tail_pct must be one of pcts");
        }
    }
    if (h->has_f) {
        if (!build_select_plan(p->path_pcts, (int)p->n_path_pcts, p->n_paths_global, p->dtype, &h->prob_f.plan))
            return fail(h, MC_ERR_INVALID, "path percentiles must be ascending and within [0, 100]");
        const size_t rows = (size_t)p->n_steps + 1;
        if (rows > 65535) return fail(h, MC_ERR_INVALID, "fan chart supports at most 65534 steps");
        MC_CUDA(h, h->states_f.reserve(rows));
        MC_CUDA(h, h->scratch_f.reserve(rows));
        MC_CUDA(h, h->fan.reserve(rows * p->n_path_pcts));
        h->prob_f.x = d_paths; h->prob_f.dtype = p->dtype; h->prob_f.rows = rows; h->prob_f.n = p->n_paths;
        h->prob_f.ld = p->paths_ld ? p->paths_ld : p->n_paths; h->prob_f.d_states = h->states_f.p;
        h->prob_f.d_scratch = h->scratch_f.p;
        setup_sampling(h->prob_f, R, true);
        attach_side(h, h->prob_f, h->side_f);
        if ((rc = upload_plan(h, h->prob_f, 1, s))) return rc;
    }
    if (p->hist_bins) {
        MC_CUDA(h, h->edges.reserve(2 * ((size_t)p->hist_bins + 2)));
        MC_CUDA(h, h->hist_total.reserve(p->hist_bins));
        h->edges_bins = 0;                   // the closing stage writes its own edges: the fused path's cache is gone
    }
    const Stage0Layout L = layout_stage0(h);
    if (!d_stage_out || cap_words < (uint64_t)mc_stage_max_words(p))
        return fail(h, MC_ERR_INVALID, "stage buffer smaller than mc_stage_max_words()");
    // The terminal problem may run beside the fan chart's when nothing of what either touches in the stage buffer can
    // meet: a pass writes [0, words_t) for the terminal row and [words_t, ...) for the fan chart, the final stage
    // (tail + histogram, terminal side) [0, 2 + bins) -- which must stay inside the terminal part.
    h->overlap_t = h->allow_overlap && !h->no_overlap && h->has_t && h->has_f && R == 1 && pass_group_of(*p) == 1 &&
                   !(p->flags & MC_FLAG_PATHS_VOLATILE) && h->prob_t.pilot_in_place && h->prob_f.pilot_in_place &&
                   final_stage_words(*p) <= pass_words(h->prob_t);
    unsigned long long* out = static_cast<unsigned long long*>(d_stage_out);

    MC_CUDA(h, cudaEventRecord(h->ev[0], s));
    SimBuffers b{};
    b.d_pool = d_pool; b.d_shocks = d_shocks; b.d_paths = d_paths; b.d_terminal = d_terminal;
    b.d_partials = h->partials.p;
    b.d_lut = h->lut.p;
    b.sink = nullptr;
    h->sim_buffers = b;
    h->has_pf = p->n_assets > 1 && pf && !(p->flags & MC_FLAG_TERMINAL_GIVEN);
    if (h->has_pf) {
        const size_t d = p->n_assets;
        h->pf_chol.assign(pf->h_chol, pf->h_chol + d * d);
        h->pf_sigma.assign(pf->h_sigma_step, pf->h_sigma_step + d);
        h->pf_drift.assign(pf->h_drift_step, pf->h_drift_step + d);
        h->pf_s0.assign(pf->h_s0, pf->h_s0 + d);
        h->pf_w.assign(pf->h_weights, pf->h_weights + d);
    }
    int grid = 0;
    if (p->n_paths && (p->flags & MC_FLAG_TERMINAL_GIVEN)) {
        MC_CUDA(h, launch_terminal_partials(d_terminal, p->dtype, p->n_paths, p->s0, h->partials.p, h->dev.sm_count, &grid, s));
        h->launches += 1;
    } else if (p->n_paths) {
        MC_CUDA(h, launch_simulation(*p, pf, b, h->dev, &grid, s));
        h->launches += 1;
    }
    MC_CUDA(h, cudaEventRecord(h->ev[1], s));
    MC_CUDA(h, launch_reduce_partials(h->partials.p, grid, out, h->has_t ? h->prob_t.m_local : 0,
                                      h->has_f ? h->prob_f.m_local : 0, s));
    h->launches += 1;
    if (h->has_t && h->prob_t.m_local) {
        MC_CUDA(h, launch_select_sample(h->prob_t, out + L.off_t, s));
        h->launches += h->prob_t.pilot_in_place ? 0 : 1;       // read in place: no sampling kernel
    }
    if (h->has_f && h->prob_f.m_local) {
        MC_CUDA(h, launch_select_sample(h->prob_f, out + L.off_f, s));
        h->launches += h->prob_f.pilot_in_place ? 0 : 1;
    }
    h->next_words = L.words;
    h->active = true;
    return MC_OK;
}

// enqueue one narrowing pass of every selection problem into `out`
static int resimulate_locked(mc_handle* h, cudaStream_t s) {
    if (h->p.n_paths == 0) return MC_OK;
    mc_portfolio pf{};
    if (h->has_pf) {
        pf.h_chol = h->pf_chol.data(); pf.h_sigma_step = h->pf_sigma.data(); pf.h_drift_step = h->pf_drift.data();
        pf.h_s0 = h->pf_s0.data(); pf.h_weights = h->pf_w.data();
    }
    int grid = 0;
    MC_CUDA(h, launch_simulation(h->p, h->has_pf ? &pf : nullptr, h->sim_buffers, h->dev, &grid, s));
    h->launches += 1;
    return MC_OK;
}

// MC_FLAG_PATHS_VOLATILE: the rows may have been overwritten since pass 1.  A later pass works from the side
// lists pass 1 kept; only if some block of it cannot (overflowed list, interval outside the brackets) is the
// run simulated again first.  Costs one small device->host read.
static int restore_rows_if_needed(mc_handle* h, cudaStream_t s) {
    if (!(h->p.flags & MC_FLAG_PATHS_VOLATILE) || !h->has_f || h->pass_idx == 0) return MC_OK;
    MC_CUDA(h, cudaMemsetAsync(h->side_blocks.p, 0, sizeof(unsigned int), s));
    MC_CUDA(h, launch_side_check(h->prob_f, h->dev.sm_count, h->side_blocks.p, s));
    MC_CUDA(h, cudaMemcpyAsync(h->h_side_blocks.p, h->side_blocks.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, s));
    MC_CUDA(h, cudaStreamSynchronize(s));
    h->launches += 1;
    if (*h->h_side_blocks.p == 0) return MC_OK;
    h->resimulations += 1;
    return resimulate_locked(h, s);
}


// mc_params.pass_group > 1: the handles of one GPU that share a job (chunks of a fan chart by regeneration) write their
// pass outputs into ONE buffer -- the leader's, which the leader clears -- so that the GPU hands the other GPUs one
// buffer per pass instead of one per chunk.
static int enqueue_pass(mc_handle* h, unsigned long long* out, cudaStream_t s, const mc_handle* leader = nullptr) {
    const PassLayout L = layout_pass(h);
    const bool first = h->pass_idx == 0;
    int rc = restore_rows_if_needed(h, s);
    if (rc) return rc;
    bool zero = true;
    if (pass_group_of(h->p) > 1 && leader) {
        if (!leader->last_pass_out) return fail(h, MC_ERR_STATE, "shared pass output: the leader has not enqueued its pass");
        out = leader->last_pass_out;
        zero = false;
    }
    h->last_pass_out = out;
    if (h->has_t) {
        MC_CUDA(h, launch_select_pass(h->prob_t, first, out + L.off_t, h->dev.sm_count, h->overlap_t ? h->side : s, zero));
        h->launches += h->prob_t.n ? 1 : 0;
    }
    if (h->has_f) {
        MC_CUDA(h, launch_select_pass(h->prob_f, first, out + L.off_f, h->dev.sm_count, s, zero));
        h->launches += h->prob_f.n ? 1 : 0;
    }
    h->pass_idx += 1;
    h->next_words = L.words;
    return MC_OK;
}

// percentiles from the decided keys, then the final stage: [tail count, tail sum, histogram counts...]
static int enqueue_closing(mc_handle* h, unsigned long long* out, cudaStream_t s) {
    const mc_params& p = h->p;
    cudaStream_t st = h->overlap_t ? h->side : s;      // everything about the terminal values
    if (h->has_t) {
        double* pct = h->pct_t.p;
        MC_CUDA(h, launch_select_finish(h->prob_t, pct, pct + MC_MAX_PCTS, pct + 2 * MC_MAX_PCTS, st));
        MC_CUDA(h, launch_publish_percentiles(pct, pct + MC_MAX_PCTS, pct + 2 * MC_MAX_PCTS, (int)p.n_pcts,
                                              h->tail_idx, h->res.p, st));
        h->launches += 2;
    }
    if (h->has_f) {
        MC_CUDA(h, launch_select_finish(h->prob_f, h->fan.p, nullptr, nullptr, s));
        h->launches += 1;
    }
    const bool want_tail = h->has_t && h->tail_idx >= 0;
    MC_CUDA(h, launch_tail_hist(h->d_terminal, p.dtype, p.n_paths, want_tail, p.hist_bins, p.hist_lo, p.hist_hi,
                                h->res.p, h->edges.p, h->tail_scratch.p, out, h->dev.sm_count, st));
    if (h->overlap_t) {       // the caller's stream takes the terminal side back in: the final merge reads both
        MC_CUDA(h, cudaEventRecord(h->ev_join, h->side));
        MC_CUDA(h, cudaStreamWaitEvent(s, h->ev_join, 0));
    }
    h->launches += (p.hist_bins ? 1 : 0) + ((p.n_paths && (want_tail || p.hist_bins)) ? 1 : 0) + (want_tail && p.n_paths ? 1 : 0);
    h->next_words = final_stage_words(p);
    h->phase = 2;
    return MC_OK;
}

// `leader`: a handle of this process and device that has just consumed the SAME gathered buffer (fan chart by
// regeneration: the chunks of one GPU are virtual ranks of one job).  Its merged selection state is what this handle
// would compute from the buffer too, so it is copied instead of computed again.
static int stage_next_locked(mc_handle* h, const void* d_gathered, uint32_t n_ranks, void* d_stage_out, int* done,
                             cudaStream_t s, const mc_handle* leader = nullptr) {
    if (!h->active) return fail(h, MC_ERR_STATE, "mc_stage_next without mc_stage_begin");
    if (!d_gathered || !d_stage_out || n_ranks == 0 || !done) return fail(h, MC_ERR_INVALID, "bad stage arguments");
    {   // a gathered pass output holds one buffer per pass GROUP, every other stage one per (virtual) rank
        const unsigned int expect = h->phase == 1 ? ranks_of(h->p) / pass_group_of(h->p) : ranks_of(h->p);
        if (n_ranks != expect) return fail(h, MC_ERR_INVALID, "n_ranks differs from what this stage gathers (mc_params.n_ranks, pass_group)");
    }
    MC_DEVICE(h);
    const unsigned long long* g = static_cast<const unsigned long long*>(d_gathered);
    unsigned long long* out = static_cast<unsigned long long*>(d_stage_out);
    const mc_params& p = h->p;
    const size_t stride = h->next_words;      // words per rank in `gathered`
    const bool selecting = h->has_t || h->has_f;
    *done = 0;

    if (h->phase == 2) {                      // gathered = final stage
        if (h->pending_unchecked) {
            // The closing stage was enqueued right behind the second narrowing pass without waiting for
            // the device's "rows still undecided" counter (identical on every rank).  It has long arrived.
            h->pending_unchecked = false;
            MC_CUDA(h, cudaEventSynchronize(h->ev_pending));
            if (*h->h_pending.p != 0) {       // a bracket missed or a heavy tie: narrow again, then close again
                if (h->pass_idx >= SEL_MAX_PASSES) return fail(h, MC_ERR_STATE, "selection did not converge");
                h->phase = 1;
                return enqueue_pass(h, out, s, leader);
            }
        }
        MC_CUDA(h, launch_merge_final(g, n_ranks, stride, p.hist_bins, h->res.p, h->hist_total.p, s));
        h->launches += 1;
        MC_CUDA(h, cudaEventRecord(h->ev[2], s));
        *done = 1;
        h->active = false;
        return MC_OK;
    }
    if (h->phase == 0) {                      // gathered = stage 0: moments + pilot samples
        MC_CUDA(h, launch_merge_moments(g, n_ranks, stride, p.s0, h->res.p, s));
        h->launches += 1;
        if (!selecting) return enqueue_closing(h, out, s);
        const Stage0Layout L = layout_stage0(h);
        if (leader) {
            if (h->has_t) MC_CUDA(h, cudaMemcpyAsync(h->states_t.p, leader->states_t.p, sizeof(RowSel), cudaMemcpyDeviceToDevice, s));
            if (h->has_f) MC_CUDA(h, cudaMemcpyAsync(h->states_f.p, leader->states_f.p, (size_t)h->prob_f.rows * sizeof(RowSel), cudaMemcpyDeviceToDevice, s));
        } else {
            if (h->has_f) {
                MC_CUDA(h, launch_select_brackets(h->prob_f, g, L.off_f, 7, n_ranks, stride, h->dev.max_smem_optin, s));
                h->launches += 1;
            }
            if (h->overlap_t) {
                // fork: from here on the terminal problem has its own stream.  Everything that reads the stage-0 header
                // on the caller's stream is enqueued above; the first thing the side stream does to the buffer is to
                // clear the terminal part of the pass layout, which lies over that header.
                MC_CUDA(h, cudaEventRecord(h->ev_fork, s));
                MC_CUDA(h, cudaStreamWaitEvent(h->side, h->ev_fork, 0));
            }
            if (h->has_t) {
                MC_CUDA(h, launch_select_brackets(h->prob_t, g, L.off_t, 6, n_ranks, stride, h->dev.max_smem_optin,
                                                  h->overlap_t ? h->side : s));
                h->launches += 1;
            }
        }
        h->phase = 1;
        return enqueue_pass(h, out, s, leader);
    }
    // gathered = output of narrowing pass `pass_idx`
    const PassLayout L = layout_pass(h);
    // the counter is read only behind the second and later scans; the first one may add to whatever is there
    if (leader) {        // the leader's scan of this very buffer: its intervals and its count of undecided rows
        if (h->has_t) MC_CUDA(h, cudaMemcpyAsync(h->states_t.p, leader->states_t.p, sizeof(RowSel), cudaMemcpyDeviceToDevice, s));
        if (h->has_f) MC_CUDA(h, cudaMemcpyAsync(h->states_f.p, leader->states_f.p, (size_t)h->prob_f.rows * sizeof(RowSel), cudaMemcpyDeviceToDevice, s));
        MC_CUDA(h, cudaMemcpyAsync(h->pending.p, leader->pending.p, sizeof(unsigned int), cudaMemcpyDeviceToDevice, s));
    } else {
        if (h->pass_idx >= 2) MC_CUDA(h, cudaMemsetAsync(h->pending.p, 0, sizeof(unsigned int), s));
        if (h->pass_idx >= 2 && h->overlap_t) MC_CUDA(h, cudaMemsetAsync(h->pending_t.p, 0, sizeof(unsigned int), h->side));
        if (h->has_t) {
            MC_CUDA(h, launch_select_scan(h->prob_t, g + L.off_t, n_ranks, stride, h->overlap_t ? h->pending_t.p : h->pending.p,
                                          h->overlap_t ? h->side : s));
            h->launches += 2;                 // per interval, then per row
        }
        if (h->has_f) {
            MC_CUDA(h, launch_select_scan(h->prob_f, g + L.off_f, n_ranks, stride, h->pending.p, s));
            h->launches += 2;
        }
    }
    if (h->pass_idx < 2) return enqueue_pass(h, out, s, leader);
    // the second pass normally decides everything: go on to the closing stage and look at the counter later
    // (closing first: with the terminal problem on its own stream it ends with that stream joining this one, and the
    // counter of that side is only final then)
    int rc = enqueue_closing(h, out, s);
    if (rc) return rc;
    u32_to_host_kernel<<<1, 1, 0, s>>>(h->h_pending.p, h->pending.p, h->overlap_t ? h->pending_t.p : nullptr);
    MC_CUDA(h, cudaGetLastError());
    h->launches += 1;
    MC_CUDA(h, cudaEventRecord(h->ev_pending, s));
    h->pending_unchecked = true;
    return MC_OK;
}

static int fetch_locked(mc_handle* h, mc_result* r, uint64_t* h_hist, double* h_fan, cudaStream_t s) {
    if (h->active) return fail(h, MC_ERR_STATE, "stages not finished");
    if (!r) return fail(h, MC_ERR_INVALID, "h_result is NULL");
    MC_DEVICE(h);
    const mc_params& p = h->p;
    const size_t fan_n = h->has_f ? ((size_t)p.n_steps + 1) * p.n_path_pcts : 0;
    const size_t bytes = sizeof(DeviceResult) + (size_t)p.hist_bins * 8 + fan_n * 8;
    MC_CUDA(h, h->h_stage.reserve(bytes));
    unsigned char* hp = h->h_stage.p;
    static_assert(sizeof(DeviceResult) % 8 == 0, "DeviceResult is copied in 8-byte words");
    const bool want_hist = p.hist_bins && h_hist, want_fan = fan_n && h_fan;
    if (bytes <= HOST_STORE_MAX_BYTES) {
        const size_t na = sizeof(DeviceResult) / 8, nb = want_hist ? (size_t)p.hist_bins : 0, nc = want_fan ? fan_n : 0;
        // the layout below keeps the histogram's slot even when it is not wanted
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(hp);
        const unsigned int blocks = (unsigned int)((na + nb + nc + 255) / 256 < 64 ? (na + nb + nc + 255) / 256 : 64);
        h->launches += (want_hist || !want_fan) ? 1 : 2;
        if (want_hist || !want_fan) {
            words_to_host_kernel<<<blocks, 256, 0, s>>>(dst, reinterpret_cast<const unsigned long long*>(h->res.p), na,
                                                        reinterpret_cast<const unsigned long long*>(h->hist_total.p), nb,
                                                        reinterpret_cast<const unsigned long long*>(h->fan.p), nc);
        } else {      // fan chart without histogram: two pieces around the unused histogram slot
            words_to_host_kernel<<<blocks, 256, 0, s>>>(dst, reinterpret_cast<const unsigned long long*>(h->res.p), na,
                                                        nullptr, 0, nullptr, 0);
            words_to_host_kernel<<<blocks, 256, 0, s>>>(dst + na + (size_t)p.hist_bins,
                                                        reinterpret_cast<const unsigned long long*>(h->fan.p), nc, nullptr, 0,
                                                        nullptr, 0);
        }
        MC_CUDA(h, cudaGetLastError());
    } else {
        MC_CUDA(h, cudaMemcpyAsync(hp, h->res.p, sizeof(DeviceResult), cudaMemcpyDeviceToHost, s));
        if (want_hist)
            MC_CUDA(h, cudaMemcpyAsync(hp + sizeof(DeviceResult), h->hist_total.p, (size_t)p.hist_bins * 8, cudaMemcpyDeviceToHost, s));
        if (want_fan)
            MC_CUDA(h, cudaMemcpyAsync(hp + sizeof(DeviceResult) + (size_t)p.hist_bins * 8, h->fan.p, fan_n * 8, cudaMemcpyDeviceToHost, s));
    }
    MC_CUDA(h, cudaStreamSynchronize(s));
    DeviceResult d;
    memcpy(&d, hp, sizeof(d));
    memset(r, 0, sizeof(*r));
    r->n = d.n; r->n_above = d.n_above; r->mean = d.mean; r->std = d.std; r->min = d.vmin; r->max = d.vmax;
    if (h->has_t) {
        memcpy(r->pct, d.pct, sizeof(d.pct));
        memcpy(r->pct_lo, d.pct_lo, sizeof(d.pct_lo));
        memcpy(r->pct_hi, d.pct_hi, sizeof(d.pct_hi));
    }
    r->tail_count = d.tail_count; r->tail_sum = d.tail_sum;
    r->hist_lo = d.hist_lo; r->hist_hi = d.hist_hi;
    if (p.hist_bins && h_hist) memcpy(h_hist, hp + sizeof(DeviceResult), (size_t)p.hist_bins * 8);
    if (fan_n && h_fan) {
        // device layout [row][pct] -> caller layout [pct][row] (path_percentiles['p5'] is a vector over time)
        const double* src = reinterpret_cast<const double*>(hp + sizeof(DeviceResult) + (size_t)p.hist_bins * 8);
        const size_t rows = (size_t)p.n_steps + 1;
        for (size_t rr = 0; rr < rows; ++rr)
            for (unsigned k = 0; k < p.n_path_pcts; ++k) h_fan[k * rows + rr] = src[rr * p.n_path_pcts + k];
    }
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]) == cudaSuccess) r->sim_ms = ms;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[2]) == cudaSuccess) r->total_ms = ms;
    r->launches = h->launches;
    return MC_OK;
}

int mc_stage_begin(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool, const void* d_shocks,
                   void* d_paths, void* d_terminal, void* d_stage_out, uint64_t cap_words, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    return stage_begin_locked(h, p, pf, d_pool, d_shocks, d_paths, d_terminal, d_stage_out, cap_words,
                              static_cast<cudaStream_t>(stream));
}

int mc_stage_next(mc_handle* h, const void* d_gathered, uint32_t n_ranks, void* d_stage_out, int* done, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    return stage_next_locked(h, d_gathered, n_ranks, d_stage_out, done, static_cast<cudaStream_t>(stream));
}

int mc_stage_adopt(mc_handle* h, mc_handle* leader, const void* d_gathered, uint32_t n_ranks, void* d_stage_out, int* done,
                   void* stream) {
    if (!h || !leader || h == leader) return MC_ERR_INVALID;
    std::scoped_lock lock(h->mu, leader->mu);
    if (leader->device != h->device) return fail(h, MC_ERR_INVALID, "mc_stage_adopt: the leader lives on another device");
    if (!h->active) return fail(h, MC_ERR_STATE, "mc_stage_adopt without mc_stage_begin");
    // the leader has taken this step already: it is one step ahead, or finished
    if (h->has_t != leader->has_t || h->has_f != leader->has_f || (h->has_f && h->prob_f.rows != leader->prob_f.rows) ||
        h->prob_t.plan.n_targets != leader->prob_t.plan.n_targets)
        return fail(h, MC_ERR_STATE, "mc_stage_adopt: the leader runs a different job");
    return stage_next_locked(h, d_gathered, n_ranks, d_stage_out, done, static_cast<cudaStream_t>(stream), leader);
}

int mc_stage_shared(mc_handle* h) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    return (h->active && h->phase == 1 && pass_group_of(h->p) > 1) ? 1 : 0;
}

int mc_stage_reads_paths(mc_handle* h) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!h->active) return fail(h, MC_ERR_STATE, "no staged run in flight");
    if (!h->has_f) return 0;
    if (h->phase == 0) return 1;                              // brackets, then the first pass
    if (h->p.flags & MC_FLAG_PATHS_VOLATILE) return 0;        // later passes: the handle restores the rows itself
    if (h->phase == 1) return h->pass_idx < 2 ? 1 : 0;        // scan, then pass 2 -- or the closing stage
    if (h->pending_unchecked) {                               // final stage, unless rows are still undecided
        MC_DEVICE(h);
        MC_CUDA(h, cudaEventSynchronize(h->ev_pending));
        return *h->h_pending.p != 0 ? 1 : 0;
    }
    return 0;
}

int mc_stage_resimulate(mc_handle* h, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!h->active) return fail(h, MC_ERR_STATE, "no staged run in flight");
    MC_DEVICE(h);
    return resimulate_locked(h, static_cast<cudaStream_t>(stream));
}

int mc_fetch_result(mc_handle* h, mc_result* r, uint64_t* h_hist, double* h_fan, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (h->fused_pending) {
        if (!r) return fail(h, MC_ERR_INVALID, "h_result is NULL");
        int status = 0;
        return fused_fetch(h, r, h_hist, static_cast<cudaStream_t>(stream), &status);
    }
    return fetch_locked(h, r, h_hist, h_fan, static_cast<cudaStream_t>(stream));
}

int mc_run_sharded(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool, const void* d_shocks,
                   void* d_paths, void* d_terminal, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!p) return fail(h, MC_ERR_INVALID, "params is NULL");
    if (!fused_eligible(*p)) return fail(h, MC_ERR_INVALID, "mc_run_sharded: this request needs the staged protocol (fan chart, data-dependent histogram range)");
    // one rank, no peers: the enqueue half of mc_run (a caller that keeps several jobs in flight on several handles)
    const bool alone = ranks_of(*p) == 1 && !(h->peers_attached && h->peers.n_ranks != 1);
    if (alone && p->n_paths != p->n_paths_global) return fail(h, MC_ERR_INVALID, "n_ranks = 1: n_paths must equal n_paths_global");
    return fused_enqueue(h, p, pf, d_pool, d_shocks, d_paths, d_terminal, !alone, static_cast<cudaStream_t>(stream));
}

int mc_peer_export(mc_handle* h, uint32_t n_ranks, void* out_ipc_handle, void** out_local_ptr, uint64_t* out_bytes) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (n_ranks < 1 || n_ranks > (uint32_t)FUSED_MAX_RANKS) return fail(h, MC_ERR_INVALID, "mc_peer_export: 1..16 ranks");
    if (h->peers_attached) return fail(h, MC_ERR_STATE, "mc_peer_export: already attached");
    MC_DEVICE(h);
    h->xchg.release();                       // a fresh allocation: its IPC handle names exactly this buffer
    h->lay = XchgLayout{};
    int rc = ensure_xchg(h, n_ranks, PEER_CAND_BYTES);
    if (rc) return rc;
    if ((rc = fused_reserve(h, FUSED_MAX_NPH_BINS))) return rc;
    h->edges_bins = 0;                       // the edge buffer may have moved
    MC_CUDA(h, cudaDeviceSynchronize());
    if (out_ipc_handle) {
        cudaIpcMemHandle_t ih;
        MC_CUDA(h, cudaIpcGetMemHandle(&ih, h->xchg.p));
        static_assert(sizeof(ih) == 64, "cudaIpcMemHandle_t is 64 bytes");
        memcpy(out_ipc_handle, &ih, sizeof(ih));
    }
    if (out_local_ptr) *out_local_ptr = h->xchg.p;
    if (out_bytes) *out_bytes = h->lay.total;
    return MC_OK;
}

int mc_peer_attach(mc_handle* h, uint32_t rank, uint32_t n_ranks, const void* ipc_handles, void* const* local_ptrs) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!h->xchg.p || h->lay.n_ranks != n_ranks || rank >= n_ranks) return fail(h, MC_ERR_STATE, "mc_peer_attach: call mc_peer_export(n_ranks) first");
    if (h->peers_attached) return fail(h, MC_ERR_STATE, "mc_peer_attach: already attached");
    if (!ipc_handles && !local_ptrs) return fail(h, MC_ERR_INVALID, "mc_peer_attach: no peer buffers given");
    MC_DEVICE(h);
    Peers P{};
    P.n_ranks = n_ranks; P.me = rank; P.seq = 0;
    for (uint32_t r = 0; r < n_ranks; ++r) {
        if (r == rank) { P.base[r] = h->xchg.p; continue; }
        if (local_ptrs) { P.base[r] = static_cast<unsigned char*>(local_ptrs[r]); continue; }
        cudaIpcMemHandle_t ih;
        memcpy(&ih, static_cast<const unsigned char*>(ipc_handles) + 64 * (size_t)r, 64);
        void* q = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&q, ih, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            for (void* o : h->ipc_opened) cudaIpcCloseMemHandle(o);
            h->ipc_opened.clear();
            return cuda_fail(h, e, "cudaIpcOpenMemHandle");
        }
        h->ipc_opened.push_back(q);
        P.base[r] = static_cast<unsigned char*>(q);
    }
    h->peers = P;
    h->peers_attached = true;
    h->fused_seq = 0;
    h->fused_dirty = true;
    return MC_OK;
}

int mc_debug_set_tuning(mc_handle* h, int side_cap, int no_side_lists, int waves, int nblock) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    h->side_cap_override = side_cap;
    h->no_side_lists = (no_side_lists & 1) != 0;
    h->no_overlap = (no_side_lists & 2) != 0;
    set_sim_tuning(waves, nblock);
    return MC_OK;
}

int mc_debug_set_pilot(uint32_t in_place_samples) {
    set_pilot_in_place_size(in_place_samples);
    return MC_OK;
}

int64_t mc_debug_counter(mc_handle* h, int which) {
    if (!h) return -1;
    std::lock_guard<std::mutex> lock(h->mu);
    switch (which) {
        case 0: return (int64_t)h->n_fused;
        case 1: return (int64_t)h->n_fused_fallback;
        case 2: return (int64_t)h->last_fused_status;
        case 3: return (int64_t)h->resimulations;
        default: return -1;
    }
}

int mc_run(mc_handle* h, const mc_params* p, const mc_portfolio* pf, const void* d_pool, const void* d_shocks,
           void* d_paths, void* d_terminal, mc_result* h_result, uint64_t* h_hist, double* h_fan, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!p) return fail(h, MC_ERR_INVALID, "params is NULL");
    if (p->struct_size != sizeof(mc_params)) return fail(h, MC_ERR_VERSION, "mc_params struct_size mismatch");
    if (p->n_paths != p->n_paths_global || p->path_offset + p->n_paths < p->n_paths || p->n_ranks > 1)
        return fail(h, MC_ERR_INVALID, "mc_run is single-rank: n_paths must equal n_paths_global");
    MC_DEVICE(h);
    if (small_eligible(*p) && small_rows_smem(p->dtype, p->n_paths) <= (size_t)h->dev.max_smem_optin - 2048 && h_result) {
        bool done = false;
        int rc = small_run(h, p, pf, d_pool, d_shocks, d_paths, d_terminal, h_result, h_fan, s, &done);
        if (rc || done) return rc;
    }
    mc_params generic = *p;
    if (fused_eligible(*p) && !(h->peers_attached && h->peers.n_ranks != 1)) {
        if (!h_result) return fail(h, MC_ERR_INVALID, "h_result is NULL");
        int rc = fused_enqueue(h, p, pf, d_pool, d_shocks, d_paths, d_terminal, false, s);
        if (rc) return rc;
        int status = 0;
        if ((rc = fused_fetch(h, h_result, h_hist, s, &status))) return rc;
        if (status == FUSED_OK) return MC_OK;
        // the window missed or a bin was too full: the terminal values are there, reduce them the general way
        generic.flags |= MC_FLAG_TERMINAL_GIVEN;
        generic.flags &= ~(uint32_t)MC_FLAG_STORE_PATHS;
    }
    p = &generic;
    const float fused_sim_ms = (generic.flags & MC_FLAG_TERMINAL_GIVEN) ? h_result->sim_ms : 0.f;
    const unsigned fused_launches = (generic.flags & MC_FLAG_TERMINAL_GIVEN) ? h->launches : 0u;
    const size_t cap = (size_t)mc_stage_max_words(p);
    MC_CUDA(h, h->stage.reserve(cap));
    h->allow_overlap = true;             // this call owns the stage buffer from the first stage to the fetch
    int rc = stage_begin_locked(h, p, pf, d_pool, d_shocks, d_paths, d_terminal, h->stage.p, cap, s);
    h->allow_overlap = false;
    if (rc) return rc;
    int done = 0;
    while (!done) {
        rc = stage_next_locked(h, h->stage.p, 1, h->stage.p, &done, s);
        if (rc) { if (h->overlap_t) cudaStreamSynchronize(h->side); return rc; }
    }
    rc = fetch_locked(h, h_result, h_hist, h_fan, s);
    if (rc == MC_OK && (generic.flags & MC_FLAG_TERMINAL_GIVEN)) {
        h_result->sim_ms = fused_sim_ms;
        h_result->launches += fused_launches;
    }
    return rc;
}

int mc_debug_shocks(mc_handle* h, const mc_params* p, const void* d_pool, void* d_out, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (!p || p->struct_size != sizeof(mc_params)) return fail(h, MC_ERR_VERSION, "mc_params mismatch");
    if (p->shock == MC_SHOCK_BOOTSTRAP && !d_pool) return fail(h, MC_ERR_INVALID, "pool missing");
    MC_DEVICE(h);
    MC_CUDA(h, launch_debug_shocks(*p, d_pool, h->lut.p, d_out, static_cast<cudaStream_t>(stream)));
    return MC_OK;
}

int mc_debug_philox(mc_handle* h, const uint32_t* d_ctr, uint32_t n, uint32_t key0, uint32_t key1, uint32_t* d_out,
                    void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    MC_DEVICE(h);
    MC_CUDA(h, launch_debug_philox(d_ctr, n, key0, key1, d_out, static_cast<cudaStream_t>(stream)));
    return MC_OK;
}

int mc_percentiles(mc_handle* h, const void* d_x, int32_t dtype, uint64_t rows, uint64_t n, uint64_t ld,
                   const double* pcts, uint32_t n_pcts, double* h_out, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!d_x || !pcts || !h_out || rows == 0 || n == 0 || n_pcts == 0 || rows > 65535 || ld < n ||
        (dtype != MC_F32 && dtype != MC_F64))
        return fail(h, MC_ERR_INVALID, "bad mc_percentiles arguments");
    if (h->active) return fail(h, MC_ERR_STATE, "a staged run is in flight on this handle");
    MC_DEVICE(h);
    MC_CUDA(h, h->fan.reserve(rows * n_pcts));
    int rc = select_rows(h, d_x, dtype, rows, n, ld, pcts, n_pcts, false, h->fan.p, nullptr, nullptr, nullptr, s);
    if (rc) return rc;
    MC_CUDA(h, cudaMemcpyAsync(h_out, h->fan.p, rows * n_pcts * sizeof(double), cudaMemcpyDeviceToHost, s));
    MC_CUDA(h, cudaStreamSynchronize(s));
    return MC_OK;
}

int mc_histogram(mc_handle* h, const void* d_x, int32_t dtype, uint64_t n, uint32_t bins, double lo, double hi,
                 uint64_t* h_counts, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if ((!d_x && n) || !h_counts || bins == 0 || bins > (1u << 16) || !(lo < hi) ||
        (dtype != MC_F32 && dtype != MC_F64))
        return fail(h, MC_ERR_INVALID, "bad mc_histogram arguments (explicit lo < hi required)");
    if (h->active) return fail(h, MC_ERR_STATE, "a staged run is in flight on this handle");
    MC_DEVICE(h);
    MC_CUDA(h, h->edges.reserve(2 * ((size_t)bins + 2)));
    h->edges_bins = 0;
    MC_CUDA(h, h->stage.reserve(2 + (size_t)bins));
    MC_CUDA(h, launch_tail_hist(d_x, dtype, n, false, bins, lo, hi, h->res.p, h->edges.p, h->tail_scratch.p,
                                h->stage.p, h->dev.sm_count, s));
    MC_CUDA(h, cudaMemcpyAsync(h_counts, h->stage.p + 2, (size_t)bins * 8, cudaMemcpyDeviceToHost, s));
    MC_CUDA(h, cudaStreamSynchronize(s));
    return MC_OK;
}

int mc_prepare_inputs(mc_handle* h, const double* d_close, const uint32_t* d_lengths, uint32_t n_series, uint32_t n_obs,
                      uint64_t close_ld, uint32_t first_row, int32_t pool_dtype, void* d_pool, uint64_t pool_ld,
                      mc_series_inputs* d_inputs, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (n_series && (!d_close || !d_pool || !d_inputs)) return fail(h, MC_ERR_INVALID, "mc_prepare_inputs: NULL buffer");
    if (pool_dtype != MC_F32 && pool_dtype != MC_F64) return fail(h, MC_ERR_INVALID, "mc_prepare_inputs: bad pool dtype");
    if (close_ld < n_obs) return fail(h, MC_ERR_INVALID, "mc_prepare_inputs: close_ld < n_obs");
    const uint32_t start = first_row > 20u ? first_row : 20u;
    if (n_obs > start && pool_ld < (uint64_t)(n_obs - start))
        return fail(h, MC_ERR_INVALID, "mc_prepare_inputs: pool_ld < n_obs - max(first_row, 20)");
    MC_DEVICE(h);
    MC_CUDA(h, launch_prepare_inputs(d_close, d_lengths, n_series, n_obs, close_ld, first_row, pool_dtype, d_pool, pool_ld,
                                     d_inputs, static_cast<cudaStream_t>(stream)));
    return MC_OK;
}

int mc_calibrate_garch(mc_handle* h, const double* d_pool, uint64_t pool_ld, const mc_series_inputs* d_inputs,
                       uint32_t n_series, uint32_t min_history, uint32_t levels, mc_garch_fit* d_fit, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    if (n_series && (!d_pool || !d_inputs || !d_fit)) return fail(h, MC_ERR_INVALID, "mc_calibrate_garch: NULL buffer");
    if (levels < 1 || levels > 16) return fail(h, MC_ERR_INVALID, "mc_calibrate_garch: levels must be 1..16");
    MC_DEVICE(h);
    MC_CUDA(h, launch_calibrate_garch(d_pool, pool_ld, d_inputs, n_series, min_history, levels, d_fit,
                                      static_cast<cudaStream_t>(stream)));
    return MC_OK;
}

int mc_run_batch(mc_handle* h, const mc_batch_params* p, const void* d_pool, uint64_t pool_ld,
                 const mc_series_inputs* d_inputs, const mc_garch_fit* d_fit, void* d_paths, void* d_terminal,
                 mc_result* h_results, double* h_fan, void* stream) {
    if (!h) return MC_ERR_INVALID;
    std::lock_guard<std::mutex> lock(h->mu);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!p) return fail(h, MC_ERR_INVALID, "params is NULL");
    if (p->struct_size != sizeof(mc_batch_params) || p->version != MC_ABI_VERSION)
        return fail(h, MC_ERR_VERSION, "mc_batch_params struct_size/version mismatch");
    if (h->active) return fail(h, MC_ERR_STATE, "a staged run is in flight on this handle");
    const bool store = p->flags & MC_FLAG_STORE_PATHS, fan = (p->flags & MC_FLAG_FAN_CHART) && p->n_path_pcts;
    if (p->n_series == 0 || p->n_series > 65535) return fail(h, MC_ERR_INVALID, "n_series must be 1..65535");
    if (p->n_paths == 0 || p->n_paths >= (1ull << 32)) return fail(h, MC_ERR_INVALID, "n_paths out of range");
    if (p->n_steps == 0 || p->n_steps > 1000000u) return fail(h, MC_ERR_INVALID, "n_steps out of range");
    if (!d_pool || !d_inputs || !d_terminal || !h_results) return fail(h, MC_ERR_INVALID, "mc_run_batch: NULL buffer");
    if (store && !d_paths) return fail(h, MC_ERR_INVALID, "d_paths is NULL");
    if (fan && (!store || !h_fan)) return fail(h, MC_ERR_INVALID, "fan chart needs stored paths and h_fan");
    if (p->n_pcts > MC_MAX_PCTS || p->n_path_pcts > MC_MAX_PATH_PCTS) return fail(h, MC_ERR_INVALID, "too many percentiles");
    int tail_idx = -1;
    if (p->tail_pct >= 0.0) {
        for (unsigned k = 0; k < p->n_pcts; ++k) if (p->pcts[k] == p->tail_pct) tail_idx = (int)k;
        if (tail_idx < 0) return fail(h, MC_ERR_INVALID, "tail_pct must be one of pcts");
    }
    MC_DEVICE(h);
    const size_t B = p->n_series, np = p->n_pcts ? p->n_pcts : 1;
    MC_CUDA(h, h->batch_pct.reserve(3 * B * np));
    MC_CUDA(h, h->batch_res.reserve(B));
    unsigned launches = 0;

    MC_CUDA(h, cudaEventRecord(h->ev[0], s));
    BatchSimArgs a{};
    a.n_paths = p->n_paths; a.ld_terminal = p->n_paths; a.ld_paths = p->n_paths; a.pool_ld = pool_ld; a.seed = p->seed;
    a.n_steps = p->n_steps; a.n_series = p->n_series; a.min_history = p->min_history;
    a.omega = p->omega; a.alpha = p->alpha; a.beta = p->beta;
    a.pool = static_cast<const double*>(d_pool); a.inputs = d_inputs; a.fit = d_fit;
    a.paths = static_cast<double*>(d_paths); a.terminal = static_cast<double*>(d_terminal);
    MC_CUDA(h, launch_batch_simulation(a, store, h->dev.sm_count, s));
    ++launches;
    MC_CUDA(h, cudaEventRecord(h->ev[1], s));

    double* d_pct = h->batch_pct.p;
    double* d_lo = d_pct + B * np;
    double* d_hi = d_lo + B * np;
    int rc;
    if (p->n_pcts) {
        if ((rc = select_rows(h, d_terminal, MC_F64, B, p->n_paths, p->n_paths, p->pcts, p->n_pcts, true, d_pct, d_lo,
                              d_hi, &launches, s)))
            return rc;
    }
    MC_CUDA(h, launch_batch_stats(a.terminal, p->n_paths, p->n_paths, p->n_series, d_inputs, d_fit, p->min_history, d_pct, d_lo,
                                  d_hi, (int)p->n_pcts, tail_idx, h->batch_res.p, s));
    ++launches;
    MC_CUDA(h, cudaMemcpyAsync(h_results, h->batch_res.p, B * sizeof(mc_result), cudaMemcpyDeviceToHost, s));
    if (fan) {
        // rows of the fan-chart problem: (history, step) pairs, at most 65535 per selection
        const size_t rows_per = (size_t)p->n_steps + 1;
        if (rows_per > 65535) return fail(h, MC_ERR_INVALID, "fan chart supports at most 65534 steps");
        const size_t series_per_call = 65535 / rows_per;
        MC_CUDA(h, h->fan.reserve(series_per_call * rows_per * p->n_path_pcts));
        for (size_t b0 = 0; b0 < B; b0 += series_per_call) {
            const size_t nb = (B - b0 < series_per_call) ? B - b0 : series_per_call;
            const double* x = a.paths + b0 * rows_per * p->n_paths;
            if ((rc = select_rows(h, x, MC_F64, nb * rows_per, p->n_paths, p->n_paths, p->path_pcts, p->n_path_pcts, true,
                                  h->fan.p, nullptr, nullptr, &launches, s)))
                return rc;
            MC_CUDA(h, cudaMemcpyAsync(h_fan + b0 * rows_per * p->n_path_pcts, h->fan.p,
                                       nb * rows_per * p->n_path_pcts * sizeof(double), cudaMemcpyDeviceToHost, s));
            if (b0 + series_per_call < B) MC_CUDA(h, cudaStreamSynchronize(s));    // h->fan is reused
        }
    }
    MC_CUDA(h, cudaEventRecord(h->ev[2], s));
    MC_CUDA(h, cudaStreamSynchronize(s));
    float sim_ms = 0.f, total_ms = 0.f;
    MC_CUDA(h, cudaEventElapsedTime(&sim_ms, h->ev[0], h->ev[1]));
    MC_CUDA(h, cudaEventElapsedTime(&total_ms, h->ev[0], h->ev[2]));
    for (size_t b = 0; b < B; ++b) {
        h_results[b].sim_ms = sim_ms; h_results[b].total_ms = total_ms;
        h_results[b].launches = launches; h_results[b].status = 0;
    }
    return MC_OK;
}

}  // extern "C"
