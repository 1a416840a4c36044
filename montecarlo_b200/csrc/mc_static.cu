// mc_static.cu -- mc_histogram_static: the fixed-bin histogram of mc_histogram for callers that must not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Nothing else in the library calls into this file.
//
// The existing entry point keeps its scratch (bin edges, per-launch counts) in buffers a handle grows with
// cudaMalloc.  Here the same scratch is ONE file-scope __device__ array of a fixed size: MC_STATIC_HIST_MAX_BINS
// bins, 84 KB, held by every GPU that loads the library until the process exits.  Consequences, all documented in
// include/mc_b200.h: sizes beyond the maximum are refused; every call shares the array, which still holds the last
// call's contents (the first kernel of a call clears what must start at zero); two calls that overlap on the GPU
// (different streams, or a captured graph replayed concurrently with itself) race on it.
//
// In HOST code the array's name is a host-side stand-in, not a device address: handing it to a kernel makes the
// kernel fault or, where the GPU can reach host memory, silently use the wrong memory.  The device address is
// asked for once per device, with that device current, by cudaGetSymbolAddress (scratch_for_device) and only that
// pointer goes to the kernels.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>

#include <mutex>

#include "mc_fused.cuh"      // np_hist_bin: np.histogram's bin of one value (shared with the fused closing)

namespace mc {
namespace {

constexpr unsigned int STATIC_MAX_BINS = MC_STATIC_HIST_MAX_BINS;
// layout of the scratch array (byte offsets, each a multiple of 256)
constexpr size_t OFF_COUNTS = 0;                                                   // u64 [MAX_BINS]       32768 B
constexpr size_t OFF_EDGES64 = OFF_COUNTS + (size_t)STATIC_MAX_BINS * 8;           // f64 [MAX_BINS + 1]   32776 B
constexpr size_t OFF_EDGES32 = (OFF_EDGES64 + ((size_t)STATIC_MAX_BINS + 1) * 8 + 255) & ~(size_t)255;   // f32 [MAX_BINS + 1]
constexpr size_t SCRATCH_BYTES = (OFF_EDGES32 + ((size_t)STATIC_MAX_BINS + 1) * 4 + 255) & ~(size_t)255;
static_assert(SCRATCH_BYTES <= 84 * 1024, "the header promises 84 KB");

static __device__ __align__(256) unsigned char g_static_hist_scratch[SCRATCH_BYTES];

// edges = np.linspace(lo, hi, bins + 1) in fp64 and, for fp32 data, rounded so that "v < edge" / "v <= hi" against
// the float64 edge is exact in fp32 (the arithmetic of hist_edges_kernel, mc_select.cu); clears the counts
__global__ void static_hist_setup_kernel(unsigned char* scratch, unsigned int bins, double lo, double hi) {
    unsigned long long* counts = reinterpret_cast<unsigned long long*>(scratch + OFF_COUNTS);
    double* edges = reinterpret_cast<double*>(scratch + OFF_EDGES64);
    float* edges_f = reinterpret_cast<float*>(scratch + OFF_EDGES32);
    const double step = __ddiv_rn(__dsub_rn(hi, lo), (double)bins);
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i <= bins; i += gridDim.x * blockDim.x) {
        const double e = (i == bins) ? hi : __dadd_rn(__dmul_rn((double)i, step), lo);
        edges[i] = e;
        float f = (float)e;
        if (i == bins) { if ((double)f > e) f = nextafterf(f, -INFINITY); }
        else if ((double)f < e) f = nextafterf(f, INFINITY);
        edges_f[i] = f;
        if (i < bins) counts[i] = 0ull;          // the array keeps the last call's counts: start at zero
    }
}

template <typename Real>
__global__ void __launch_bounds__(256) static_hist_kernel(const Real* __restrict__ x, unsigned long long n,
                                                          unsigned char* scratch, unsigned int bins) {
    __shared__ unsigned int s_counts[STATIC_MAX_BINS];
    unsigned long long* counts = reinterpret_cast<unsigned long long*>(scratch + OFF_COUNTS);
    const Real* edge = sizeof(Real) == 8 ? reinterpret_cast<const Real*>(scratch + OFF_EDGES64)
                                         : reinterpret_cast<const Real*>(scratch + OFF_EDGES32);
    for (unsigned int i = threadIdx.x; i < bins; i += blockDim.x) s_counts[i] = 0u;
    __syncthreads();
    const Real lo = edge[0], top = edge[bins], inv_w = (Real)bins / (top - lo);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int b = np_hist_bin<Real>(x[i], edge, (int)bins, lo, top, inv_w);
        if (b >= 0) atomicAdd(&s_counts[b], 1u);
    }
    __syncthreads();
    for (unsigned int i = threadIdx.x; i < bins; i += blockDim.x) {
        const unsigned int c = s_counts[i];
        if (c) atomicAdd(&counts[i], (unsigned long long)c);
    }
}

__global__ void static_hist_publish_kernel(const unsigned char* scratch, unsigned int bins, unsigned long long* out) {
    const unsigned long long* counts = reinterpret_cast<const unsigned long long*>(scratch + OFF_COUNTS);
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < bins; i += gridDim.x * blockDim.x) out[i] = counts[i];
}

// device address of the scratch array on `device`, asked for once per device with that device current
constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_ptr[MAX_DEVICES] = {};
int g_sm_count[MAX_DEVICES] = {};

cudaError_t scratch_for_device(int device, unsigned char** ptr, int* sm_count) {
    std::lock_guard<std::mutex> lock(g_setup_mu);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_hist_scratch);      // NOT &g_static_hist_scratch: see the header comment
        if (e != cudaSuccess) return e;
        int sms = 0;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) return e;
        g_scratch_ptr[device] = static_cast<unsigned char*>(p);
        g_sm_count[device] = sms;
    }
    *ptr = g_scratch_ptr[device];
    *sm_count = g_sm_count[device];
    return cudaSuccess;
}

int refuse(char* err, uint32_t cap, int code, const char* msg) {
    if (err && cap) snprintf(err, cap, "%s", msg);
    return code;
}
int refuse_cuda(char* err, uint32_t cap, cudaError_t e, const char* where) {
    if (err && cap) snprintf(err, cap, "mc_histogram_static: %s: %s", where, cudaGetErrorString(e));
    cudaGetLastError();          // reported here: it must not resurface behind a later, unrelated launch
    return MC_ERR_CUDA;
}

}  // namespace
}  // namespace mc

extern "C" int mc_histogram_static(int device_ordinal, const void* d_x, int32_t dtype, uint64_t n, uint32_t bins, double lo,
                                   double hi, uint64_t* d_counts, void* stream, char* err, uint32_t err_cap) {
    using namespace mc;
    if (err && err_cap) err[0] = '\0';
    if (bins == 0 || bins > STATIC_MAX_BINS) {
        if (err && err_cap)
            snprintf(err, err_cap, "mc_histogram_static: %u bins requested, the static scratch array holds at most %u "
                                   "(MC_STATIC_HIST_MAX_BINS); use mc_histogram", bins, STATIC_MAX_BINS);
        return MC_ERR_INVALID;
    }
    if ((!d_x && n) || !d_counts) return refuse(err, err_cap, MC_ERR_INVALID, "mc_histogram_static: NULL buffer");
    if (!(lo < hi)) return refuse(err, err_cap, MC_ERR_INVALID, "mc_histogram_static: an explicit range lo < hi is required");
    if (dtype != MC_F32 && dtype != MC_F64) return refuse(err, err_cap, MC_ERR_INVALID, "mc_histogram_static: bad dtype");
    if (device_ordinal < 0 || device_ordinal >= MAX_DEVICES) return refuse(err, err_cap, MC_ERR_INVALID, "mc_histogram_static: no such CUDA device");
    // work on the requested device, leave the calling thread's current device as it was
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return refuse_cuda(err, err_cap, e, "cudaGetDevice");
    if (prev != device_ordinal && (e = cudaSetDevice(device_ordinal)) != cudaSuccess) return refuse_cuda(err, err_cap, e, "cudaSetDevice");
    int rc = MC_OK;
    unsigned char* scratch = nullptr;
    int sms = 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if ((e = scratch_for_device(device_ordinal, &scratch, &sms)) != cudaSuccess) {
        rc = refuse_cuda(err, err_cap, e, "cudaGetSymbolAddress");
    } else {
        static_hist_setup_kernel<<<(bins + 256) / 256, 256, 0, s>>>(scratch, bins, lo, hi);
        if (n) {
            unsigned long long blocks = (n + 256ull * 8 - 1) / (256ull * 8);
            if (blocks > (unsigned long long)sms * 4) blocks = (unsigned long long)sms * 4;
            if (blocks < 1) blocks = 1;
            if (dtype == MC_F64) static_hist_kernel<double><<<(unsigned)blocks, 256, 0, s>>>(static_cast<const double*>(d_x), n, scratch, bins);
            else static_hist_kernel<float><<<(unsigned)blocks, 256, 0, s>>>(static_cast<const float*>(d_x), n, scratch, bins);
        }
        static_hist_publish_kernel<<<(bins + 255) / 256, 256, 0, s>>>(scratch, bins, reinterpret_cast<unsigned long long*>(d_counts));
        if ((e = cudaGetLastError()) != cudaSuccess) rc = refuse_cuda(err, err_cap, e, "kernel launch");
    }
    if (prev != device_ordinal && prev >= 0) cudaSetDevice(prev);
    return rc;
}
