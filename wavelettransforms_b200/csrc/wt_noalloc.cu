// wt_noalloc.cu -- wt_select_abs_noalloc_f32: the exact order-statistic select of wt_select_abs_f32 (wt_blocks.cu) for
// callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  It has no counterpart in the reference.
//
// The scratch the select kernel needs -- the job descriptor, histograms and per-job state, the candidate-key list --
// lives in ONE file-scope device array of a fixed size, so the entry point calls no cudaMalloc / cudaFree and no
// allocating library routine.  Every GPU that loads the library holds the array (WT_NOALLOC_SCRATCH_BYTES, about 4.2 MB)
// until the process exits.  All calls share it and it keeps the last call's contents: the set-up kernel below clears
// what must start at zero, and calls that overlap on the GPU (two streams, two host threads) race -- serialise them.
#include <string.h>

#include "wt_common.cuh"

#define WT_NOALLOC_MAX_N ((long long)WT_SELECT_NOALLOC_MAX_N)  // include/wtprune.h: 4,194,304 values per call

namespace {

// layout of the scratch array (every block 256-byte aligned)
constexpr size_t kOffJob = 0;                                          // WtSelJob
constexpr size_t kOffThr = 512;                                        // float thresholds[1]
constexpr size_t kOffCnt = 768;                                        // long long counts[4]
constexpr size_t kOffSel = 1024;                                       // select scratch: histograms, state, flags
constexpr size_t kSelBytes = ((size_t)WT_SEL_BINS1 * 8 + (size_t)WT_SEL_CBINS1 * 8 + 2 * (size_t)WT_SEL_CBINS * 8 +
                              (size_t)WT_SEL_COARSE * 8 + sizeof(WtSelState) + 256 + 255) / 256 * 256;
constexpr size_t kOffCand = kOffSel + kSelBytes;                       // candidate keys
constexpr size_t kCandKeys = (size_t)(WT_NOALLOC_MAX_N / 4);           // wt_select_cand_cap(n) <= n / 4 for n >= 16 Ki
constexpr size_t kScratchBytes = kOffCand + kCandKeys * sizeof(unsigned int);
static_assert(sizeof(WtSelJob) <= kOffThr, "job descriptor slot");

}  // namespace

#define WT_NOALLOC_SCRATCH_BYTES kScratchBytes

// In host code this name is a host-side stand-in, NOT a device address: the host gets the address once per device with
// cudaGetSymbolAddress (scratch_for_current_device below) and hands that pointer to the kernels.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

namespace {

// stores the job descriptor and clears everything the select expects to be zero (thresholds, counts, histograms, state,
// flags); the candidate list needs no clearing (the kernel only reads what it appended)
__global__ void __launch_bounds__(256) wt_noalloc_setup_kernel(WtSelJob job, unsigned char* scratch, unsigned int zero_bytes) {
    uint4* z = reinterpret_cast<uint4*>(scratch + kOffThr);
    const unsigned int n16 = zero_bytes >> 4;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
    if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<WtSelJob*>(scratch + kOffJob) = job;
}

// device address of the scratch array on the current device (looked up once per device, with that device current)
int scratch_for_current_device(unsigned char** out) {
    static unsigned char* ptrs[64] = {nullptr};  // racing threads store the same value
    int dev = 0;
    WT_CUDA(cudaGetDevice(&dev));
    WT_CHECK(dev >= 0 && dev < 64, WT_E_UNSUPPORTED, "device ordinal %d out of range", dev);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        WT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        WT_CHECK(p != nullptr, WT_E_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    *out = ptrs[dev];
    return WT_OK;
}

}  // namespace

extern "C" int wt_select_abs_noalloc_f32(const float* x, int64_t n, int64_t implicit_zeros, int64_t k0, int64_t k1, float* ab_out,
                                         void* stream_) {
    WT_CHECK(n >= 0 && implicit_zeros >= 0 && n + implicit_zeros >= 1, WT_E_INVALID, "empty population");
    WT_CHECK(ab_out != nullptr && (n == 0 || x != nullptr), WT_E_INVALID, "NULL argument");
    WT_CHECK(k0 >= 0 && k1 >= k0 && k1 < n + implicit_zeros, WT_E_INVALID, "ranks out of range");
    WT_CHECK(n <= WT_NOALLOC_MAX_N, WT_E_INVALID,
             "wt_select_abs_noalloc_f32: n = %lld exceeds the fixed scratch of this entry point (at most %lld values; use "
             "wt_select_abs_f32, which allocates)", (long long)n, (long long)WT_NOALLOC_MAX_N);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n == 0) {
        WT_CUDA(cudaMemsetAsync(ab_out, 0, 2 * sizeof(float), stream));
        return WT_OK;
    }
    unsigned char* scratch = nullptr;
    int rc = scratch_for_current_device(&scratch);
    if (rc) return rc;
    const unsigned int cap = wt_select_cand_cap(n);
    WT_CHECK((size_t)cap <= kCandKeys && wt_select_scratch_bytes(1) <= kSelBytes, WT_E_CUDA,
             "internal error: the select's scratch needs have outgrown the fixed array");
    WtSelJob j;
    memset(&j, 0, sizeof(j));
    j.x = x; j.out = nullptr; j.n = n; j.zeros = implicit_zeros; j.k0 = k0; j.k1 = k1; j.gamma = 0.0;
    j.lerp_mode = WT_LERP_NEP50; j.tile_begin = 0;
    j.cand = reinterpret_cast<unsigned int*>(scratch + kOffCand);
    j.cand_cap = cap;
    j.sample_shift = wt_select_sample_shift(n);
    j.n_tiles = (int)((n + WT_SEL_TILE - 1) / WT_SEL_TILE);
    j.layer = 0;
    wt_noalloc_setup_kernel<<<16, 256, 0, stream>>>(j, scratch, (unsigned int)(kOffCand - kOffThr));
    WT_CUDA(cudaGetLastError());
    WtSelArgs a;
    memset(&a, 0, sizeof(a));
    wt_select_bind(&a, scratch + kOffSel, 1);
    a.jobs = reinterpret_cast<WtSelJob*>(scratch + kOffJob);
    a.n_jobs = 1; a.n_tiles = j.n_tiles; a.any_apply = 0;
    a.thresholds = reinterpret_cast<float*>(scratch + kOffThr);
    a.counts = reinterpret_cast<long long*>(scratch + kOffCnt);
    rc = wt_select_launch(a, stream);
    if (rc) return rc;
    WT_CUDA(cudaMemcpyAsync(ab_out, &a.state->a, 2 * sizeof(float), cudaMemcpyDeviceToDevice, stream));
    return WT_OK;
}
