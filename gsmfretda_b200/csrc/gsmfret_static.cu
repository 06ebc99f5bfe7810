// gsmfret_static.cu — gsm_observed_hist_static(): the observed SgDivSr histogram for callers
// that are not allowed to allocate device memory.
//
// Who that is: a node embedded in a process whose device memory belongs to somebody else —
// a host application that has handed the whole of HBM to its own pool (a caching allocator,
// an arena sized at start-up) so that a cudaMalloc from a plug-in fails or fragments it; a
// latency-bound loop (a GA generation, a real-time acquisition callback) that cannot take the
// device-wide synchronisation cudaMalloc / cudaFree imply; or a sandboxed worker that is given
// buffers and a stream and nothing else.  Such a caller brings the input and the output as
// device pointers out of its own pool.  What it cannot bring is the kernel's working set — the
// 64-bit bin counters the blocks add into before the bins become floats — so that lives here,
// in a file-scope device array that every GPU loading libgsmfret.so carries until the process
// exits (kStaticScratchBytes, about 32 KB).
//
// The rest of the library (gsmfret_c.cu) keeps allocating what it needs at set-up; nothing in
// this file is on any of its paths.  No counterpart in the reference, which allocates freely
// (/root/reference/src/mc.cu:456-521).
#include "../../include/gsmfret_c.h"
#include "pda_device.cuh"

#include <cuda_runtime.h>
#include <cstdio>
#include <mutex>

// defined in gsmfret_c.cu: records the text gsm_last_error() returns, hands `code` back
__attribute__((visibility("hidden"))) int gsm_detail_fail(int code, const char* text);

namespace {

constexpr int kMaxBins = GSM_STATIC_MAX_BINS;
constexpr int kMaxDevices = 64;
constexpr size_t kStaticScratchBytes = (size_t)(kMaxBins + 2) * sizeof(unsigned long long);

// The working histogram: n_bins inner counters, then overflow, then underflow (boost regular
// axis, as hist_index orders them).  In HOST code this name is only a handle for the runtime's
// symbol table — it is not an address a kernel can use — so the host side below asks the
// runtime where the array lives on the current device and passes that pointer on.
static __device__ __align__(256) unsigned long long g_obs_scratch[kMaxBins + 2];

static_assert(sizeof(g_obs_scratch) == kStaticScratchBytes && kStaticScratchBytes == 32784, "the size include/gsmfret_c.h states");

// one resolved address per device ordinal, filled by the first call made with that device current
std::mutex g_setup_lock;
unsigned long long* g_scratch_on[kMaxDevices] = {};

// same binning as pda_hist_f32_kernel (pda_kernels.cu), counters in the array handed in
__global__ void obs_hist_static_kernel(const float* __restrict__ x, int64_t n, int nb,
                                       unsigned long long* __restrict__ work)
{
    extern __shared__ unsigned int sh[];
    const int nbp = nb + 2;
    for (int i = threadIdx.x; i < nbp; i += blockDim.x) sh[i] = 0u;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&sh[pda::hist_index(x[i], nb)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < nbp; i += blockDim.x)
        if (sh[i]) atomicAdd(work + i, (unsigned long long)sh[i]);
}

// counters -> floats for the caller; every counter (incl. the two outer ones) is left at zero
__global__ void obs_hist_static_finish_kernel(unsigned long long* __restrict__ work, float* __restrict__ out,
                                              int nb)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nb) out[i] = __ull2float_rn(work[i]);
    if (i < nb + 2) work[i] = 0ull;
}

int cuda_fail(const char* what, cudaError_t e)
{
    char buf[256];
    snprintf(buf, sizeof(buf), "gsm_observed_hist_static: %s returned %s(%d)", what, cudaGetErrorString(e), (int)e);
    return gsm_detail_fail(GSM_ECUDA, buf);
}

// Device address of g_obs_scratch on the CURRENT device; resolved once per device.
int scratch_of_current_device(unsigned long long** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return gsm_detail_fail(GSM_ENODEVICE, "gsm_observed_hist_static: no CUDA device: this library has no CPU path");
    }
    if (dev < 0 || dev >= kMaxDevices) return gsm_detail_fail(GSM_EINVAL, "gsm_observed_hist_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> hold(g_setup_lock);
    if (!g_scratch_on[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_obs_scratch);
        if (e != cudaSuccess) return cuda_fail("cudaGetSymbolAddress(g_obs_scratch)", e);
        if (!p) return gsm_detail_fail(GSM_ECUDA, "gsm_observed_hist_static: the scratch array has no address on this device");
        g_scratch_on[dev] = (unsigned long long*)p;
    }
    *out = g_scratch_on[dev];
    return GSM_OK;
}

}  // namespace

extern "C" int gsm_observed_hist_static(const float* d_SgDivSr, int64_t n_bursts, int n_bins, float* d_out,
                                        void* stream)
{
    if (n_bins < 1) return gsm_detail_fail(GSM_EINVAL, "gsm_observed_hist_static: n_bins must be at least 1");
    if (n_bins > kMaxBins) {
        char buf[160];
        snprintf(buf, sizeof(buf), "gsm_observed_hist_static: n_bins %d exceeds GSM_STATIC_MAX_BINS (%d), the size of the "
                 "built-in scratch; use gsm_init_data + gsm_observed_hist", n_bins, kMaxBins);
        return gsm_detail_fail(GSM_ERANGE, buf);
    }
    if (n_bursts < 0) return gsm_detail_fail(GSM_EINVAL, "gsm_observed_hist_static: negative n_bursts");
    if (!d_out || (n_bursts > 0 && !d_SgDivSr)) return gsm_detail_fail(GSM_EINVAL, "gsm_observed_hist_static: null pointer");

    unsigned long long* work = nullptr;
    const int rc = scratch_of_current_device(&work);
    if (rc != GSM_OK) return rc;

    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    // The array is shared by every call and remembers the previous one.  The finish kernel leaves it
    // cleared, but a call that failed half-way would not have: start from zero regardless.
    e = cudaMemsetAsync(work, 0, (size_t)(n_bins + 2) * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return cuda_fail("cudaMemsetAsync(scratch)", e);
    if (n_bursts > 0) {
        int64_t blocks = (n_bursts + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        obs_hist_static_kernel<<<(unsigned)blocks, 256, (size_t)(n_bins + 2) * sizeof(unsigned int), s>>>(
            d_SgDivSr, n_bursts, n_bins, work);
        e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail("obs_hist_static_kernel launch", e);
    }
    obs_hist_static_finish_kernel<<<(n_bins + 2 + 255) / 256, 256, 0, s>>>(work, d_out, n_bins);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail("obs_hist_static_finish_kernel launch", e);
    return GSM_OK;
}
