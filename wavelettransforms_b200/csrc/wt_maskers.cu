// wt_maskers.cu -- the two sibling maskers of the DWT prune path.
//
//   wt_min_weight_prune_f32 : percentage_min_pruning, /root/reference/ResNet/min_weight_pruning.py:66-74
//                             (k = int(n*pct) smallest |w| zeroed in place) + count_nonzero (:108-109)
//   wt_scatter_zero_f32     : flat[randperm(n)[:k]] = 0 + count_nonzero, random_pruning.py:53-58
//   wt_count_f32            : torch.count_nonzero and calculate_sparsity's |p| < 1e-6 count
//                             (testing_suite/eval_model.py:7-20)
//
// Min-weight is ONE launch of the select kernel of wt_select.cu in the common case: ranks k-1 and k with an interpolation
// weight of 1 make its fused level-0 apply zero every |w| < sorted[k], which are exactly the k smallest whenever
// a = sorted[k-1] < b = sorted[k].  Only when a == b (a tie across the boundary) the three small kernels below run for
// that tensor: they zero the first (k - #less) entries equal to a in flat-index order (torch.topk leaves tie order
// unspecified; lowest index first is this library's documented rule).  Otherwise they return at once.
#include <algorithm>

#include "wt_common.cuh"

namespace {

constexpr int MW_THREADS = 512;  // one WT_SEL_TILE (4096 elements) per CTA, 8 consecutive per thread

struct MwJob {
    float* w;
    long long n, k;
    int tile_begin, n_tiles;
    int sel_job;  // index into the select state
    int pad;
};

// does tensor `job` need the tie passes?  (k = 0: nothing to zero; a = 0: the ties are zeros already)
__device__ __forceinline__ bool mw_tie(const MwJob& job, const WtSelState& st) {
    return job.k > 0 && st.a == st.b && st.a > 0.0f;
}

__device__ __forceinline__ int find_mw_job(const MwJob* __restrict__ jobs, int n_jobs, int tile) {
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (__ldg(&jobs[mid].tile_begin) <= tile) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// pass A: per tile number of entries equal to a; per job number of entries below a
__global__ void __launch_bounds__(MW_THREADS)
mw_count_kernel(const MwJob* __restrict__ jobs, int n_jobs, const WtSelState* __restrict__ state,
                unsigned int* __restrict__ tile_eq, unsigned long long* __restrict__ job_less) {
    __shared__ unsigned int s[2 * (MW_THREADS / 32)];
    const int tile = blockIdx.x;
    const int j = find_mw_job(jobs, n_jobs, tile);
    const MwJob job = jobs[j];
    const float a = state[job.sel_job].a;
    if (!mw_tie(job, state[job.sel_job])) return;   // the select's own apply already zeroed exactly k entries
    const long long base = (long long)(tile - job.tile_begin) * WT_SEL_TILE;
    const int cnt = (int)min((long long)WT_SEL_TILE, job.n - base);
    unsigned int eq = 0, less = 0;
    for (int i = threadIdx.x; i < cnt; i += MW_THREADS) {
        const float v = fabsf(job.w[base + i]);
        eq += (v == a);
        less += (v < a);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        eq += __shfl_down_sync(0xffffffffu, eq, d);
        less += __shfl_down_sync(0xffffffffu, less, d);
    }
    if ((threadIdx.x & 31) == 0) { s[threadIdx.x >> 5] = eq; s[MW_THREADS / 32 + (threadIdx.x >> 5)] = less; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int e = 0, l = 0;
        for (int w = 0; w < MW_THREADS / 32; ++w) { e += s[w]; l += s[MW_THREADS / 32 + w]; }
        tile_eq[tile] = e;
        if (l) atomicAdd(&job_less[j], (unsigned long long)l);
    }
}

// pass B: exclusive scan of tile_eq within each job (one CTA per job; ties are rare, tiles few)
__global__ void __launch_bounds__(256)
mw_scan_kernel(const MwJob* __restrict__ jobs, const WtSelState* __restrict__ state, const unsigned int* __restrict__ tile_eq,
               unsigned long long* __restrict__ tile_eq_prefix) {
    __shared__ unsigned long long s_warp[8];
    __shared__ unsigned long long s_carry;
    const MwJob job = jobs[blockIdx.x];
    if (!mw_tie(job, state[job.sel_job])) return;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int t0 = 0; t0 < job.n_tiles; t0 += 256) {
        const int t = t0 + threadIdx.x;
        unsigned long long v = (t < job.n_tiles) ? tile_eq[job.tile_begin + t] : 0ull, inc = v;
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        unsigned long long woff = 0;
        for (int w = 0; w < wid; ++w) woff += s_warp[w];
        const unsigned long long carry = s_carry;
        if (t < job.n_tiles) tile_eq_prefix[job.tile_begin + t] = carry + woff + inc - v;
        __syncthreads();
        if (threadIdx.x == 255) s_carry = carry + woff + inc;
        __syncthreads();
    }
}

// pass C: zero and count
__global__ void __launch_bounds__(MW_THREADS)
mw_apply_kernel(const MwJob* __restrict__ jobs, int n_jobs, const WtSelState* __restrict__ state,
                const unsigned long long* __restrict__ tile_eq_prefix, const unsigned long long* __restrict__ job_less,
                unsigned long long* __restrict__ nonzero_out) {
    __shared__ unsigned int s_warp[MW_THREADS / 32];
    __shared__ unsigned int s_nz[MW_THREADS / 32];
    const int tile = blockIdx.x;
    const int j = find_mw_job(jobs, n_jobs, tile);
    const MwJob job = jobs[j];
    const float a = state[job.sel_job].a;
    if (!mw_tie(job, state[job.sel_job])) return;
    const long long base = (long long)(tile - job.tile_begin) * WT_SEL_TILE;
    const int cnt = (int)min((long long)WT_SEL_TILE, job.n - base);
    const bool none = (job.k <= 0), all = (job.k >= job.n);
    // number of entries equal to `a` that must still be zeroed, counted from the start of the tensor
    const long long quota = (none || all) ? 0 : job.k - (long long)job_less[j];
    const long long before = (long long)tile_eq_prefix[tile];
    // 8 consecutive elements per thread so that ranks follow the flat index
    const int i0 = threadIdx.x * 8;
    float v[8];
    unsigned int eq = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        v[i] = (i0 + i < cnt) ? job.w[base + i0 + i] : 1.0f;
        eq += (i0 + i < cnt) && !none && !all && (fabsf(v[i]) == a);
    }
    unsigned int inc = eq;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    unsigned int woff = 0;
    for (int w = 0; w < wid; ++w) woff += s_warp[w];
    long long rank = before + woff + inc - eq;  // rank of this thread's first equal element
    unsigned int nz = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (i0 + i >= cnt) break;
        const float m = fabsf(v[i]);
        bool zero = all;
        if (!none && !all) {
            if (m < a) zero = true;
            else if (m == a) { zero = (rank < quota); ++rank; }
        }
        if (zero) { job.w[base + i0 + i] = 0.0f; }
        else nz += (v[i] != 0.0f);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) nz += __shfl_down_sync(0xffffffffu, nz, d);
    if (lane == 0) s_nz[wid] = nz;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < MW_THREADS / 32; ++w) t += s_nz[w];
        if (t) atomicAdd(&nonzero_out[j], t);
    }
}

// count_nonzero per tensor: from the tie passes where they ran, else from the select's apply
__global__ void mw_finalize_kernel(const MwJob* __restrict__ jobs, int n_jobs, const WtSelState* __restrict__ state,
                                   const long long* __restrict__ counts, unsigned long long* __restrict__ nonzero_out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_jobs) return;
    if (!mw_tie(jobs[j], state[jobs[j].sel_job])) nonzero_out[j] = (unsigned long long)counts[(size_t)jobs[j].sel_job * 4 + 1];
}

__global__ void __launch_bounds__(256) scatter_zero_kernel(float* __restrict__ w, long long n,
                                                           const long long* __restrict__ idx, long long k) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < k; i += (long long)gridDim.x * blockDim.x) {
        const long long t = idx[i];
        if (t >= 0 && t < n) w[t] = 0.0f;
    }
}

__global__ void __launch_bounds__(256) count_kernel(const float* __restrict__ x, long long n, float eps,
                                                    unsigned long long* out2, int want_eps) {
    unsigned int nz = 0, be = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = x[i];
        nz += (v != 0.0f);
        be += (fabsf(v) < eps);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        nz += __shfl_down_sync(0xffffffffu, nz, d);
        be += __shfl_down_sync(0xffffffffu, be, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (nz) atomicAdd(&out2[0], (unsigned long long)nz);
        if (want_eps && be) atomicAdd(&out2[1], (unsigned long long)be);
    }
}

// grow-only device workspace per host thread and device: percentage_min_pruning is called once per model from
// its own thread (main_pruning.py:191-209), so the entry point neither allocates nor frees on the hot path
struct Workspace {
    char* ptr = nullptr;
    size_t bytes = 0;
    int device = -1;
    ~Workspace() { if (ptr) cudaFree(ptr); }  // at thread exit; an error after context teardown is harmless
};
thread_local Workspace g_ws;

int workspace(size_t bytes, char** out) {
    int dev = 0;
    WT_CUDA(cudaGetDevice(&dev));
    if (g_ws.ptr == nullptr || g_ws.bytes < bytes || g_ws.device != dev) {
        if (g_ws.ptr) { cudaFree(g_ws.ptr); g_ws.ptr = nullptr; g_ws.bytes = 0; }
        const size_t want = bytes + bytes / 4;
        cudaError_t e = cudaMalloc((void**)&g_ws.ptr, want);
        if (e != cudaSuccess) {
            g_ws.ptr = nullptr;
            wt_set_error("cudaMalloc(%zu) failed: %s (libwtprune has no CPU fallback)", want, cudaGetErrorString(e));
            return WT_E_CUDA;
        }
        g_ws.bytes = want;
        g_ws.device = dev;
    }
    *out = g_ws.ptr;
    return WT_OK;
}

int grid_for(long long n, int threads, int per_sm) {
    WtDeviceInfo di;
    if (wt_device_info(&di)) return 1;
    long long b = (n + threads - 1) / threads;
    b = std::min(b, (long long)di.sm_count * per_sm);
    return (int)std::max(b, 1LL);
}

}  // namespace

extern "C" int wt_min_weight_prune_f32(float* const* w, const int64_t* n, const int64_t* k, int n_tensors, void* stream_,
                                       int64_t* nonzero_out) {
    WT_CHECK(n_tensors >= 0, WT_E_INVALID, "n_tensors < 0");
    WT_CHECK(n_tensors == 0 || (w && n && k && nonzero_out), WT_E_INVALID, "NULL argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::vector<MwJob> jobs;
    std::vector<WtSelJob> sjobs;
    std::vector<int> owner;
    int tiles = 0;
    for (int i = 0; i < n_tensors; ++i) {
        nonzero_out[i] = 0;
        WT_CHECK(n[i] >= 0 && k[i] >= 0, WT_E_INVALID, "tensor %d: negative size or prune count", i);
        if (n[i] == 0) continue;
        WT_CHECK(w[i] != nullptr, WT_E_INVALID, "tensor %d: NULL pointer", i);
        // torch.topk raises when k > n ("selected index k out of range")
        WT_CHECK(k[i] <= n[i], WT_E_INVALID, "selected index k out of range");
        const long long nt = (n[i] + WT_SEL_TILE - 1) / WT_SEL_TILE;
        WT_CHECK(tiles + nt < 2147483647LL, WT_E_UNSUPPORTED, "too many tiles");
        MwJob m;
        m.w = w[i]; m.n = n[i]; m.k = k[i]; m.tile_begin = tiles; m.n_tiles = (int)nt; m.sel_job = (int)sjobs.size(); m.pad = 0;
        WtSelJob s;
        memset(&s, 0, sizeof(s));
        // ranks k-1 and k, interpolation weight 1: threshold = sorted[k], the apply zeroes |w| < sorted[k] in place
        // (k = 0: threshold = the minimum, nothing is below it)
        s.x = w[i]; s.out = w[i]; s.n = n[i]; s.zeros = 0;
        s.k0 = std::min<long long>(std::max<long long>(k[i] - 1, 0), n[i] - 1);
        s.k1 = std::min<long long>(k[i], n[i] - 1);
        s.gamma = k[i] > 0 ? 1.0 : 0.0; s.lerp_mode = WT_LERP_NEP50; s.tile_begin = tiles; s.n_tiles = (int)nt; s.layer = (int)sjobs.size();
        s.cand_cap = wt_select_cand_cap(n[i]);
        s.sample_shift = wt_select_sample_shift(n[i]);
        tiles += (int)nt;
        jobs.push_back(m);
        sjobs.push_back(s);
        owner.push_back(i);
    }
    const int nj = (int)jobs.size();
    if (nj == 0) return WT_OK;
    // device block: [sel jobs | mw jobs | thresholds | ZERO: select scratch | tile_eq | tile_eq_prefix | job_less | nonzero]
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    size_t off = 0;
    const size_t o_sj = off; off = al(off + nj * sizeof(WtSelJob));
    const size_t o_mj = off; off = al(off + nj * sizeof(MwJob));
    const size_t o_thr = off; off = al(off + nj * sizeof(float));
    const size_t o_zero = off;
    const size_t o_scr = off; off = al(off + wt_select_scratch_bytes(nj));
    const size_t o_cnt = off; off = al(off + (size_t)nj * 4 * sizeof(long long));
    const size_t o_teq = off; off = al(off + (size_t)tiles * sizeof(unsigned int));
    const size_t o_tpx = off; off = al(off + (size_t)tiles * sizeof(unsigned long long));
    const size_t o_less = off; off = al(off + (size_t)nj * sizeof(unsigned long long));
    const size_t o_nz = off; off = al(off + (size_t)nj * sizeof(unsigned long long));
    const size_t zero_end = off;
    for (int j = 0; j < nj; ++j) {  // candidate lists of the select fast path (not zeroed)
        const size_t o = off;
        off = al(off + (size_t)sjobs[j].cand_cap * sizeof(unsigned int));
        sjobs[j].cand = reinterpret_cast<unsigned int*>(o);  // offset for now, rebased below
    }
    char* dev = nullptr;
    int rc = workspace(off, &dev);
    if (rc) return rc;
    for (int j = 0; j < nj; ++j) sjobs[j].cand = reinterpret_cast<unsigned int*>(dev + reinterpret_cast<size_t>(sjobs[j].cand));
    std::vector<unsigned long long> nz(nj);
    do {
        cudaError_t e;
#define MW_TRY(x) if ((e = (x)) != cudaSuccess) { wt_set_error("%s failed: %s", #x, cudaGetErrorString(e)); rc = WT_E_CUDA; break; }
        MW_TRY(cudaMemcpyAsync(dev + o_sj, sjobs.data(), nj * sizeof(WtSelJob), cudaMemcpyHostToDevice, stream));
        MW_TRY(cudaMemcpyAsync(dev + o_mj, jobs.data(), nj * sizeof(MwJob), cudaMemcpyHostToDevice, stream));
        MW_TRY(cudaMemsetAsync(dev + o_zero, 0, zero_end - o_zero, stream));
        WtSelArgs a;
        memset(&a, 0, sizeof(a));
        wt_select_bind(&a, dev + o_scr, nj);
        a.jobs = reinterpret_cast<WtSelJob*>(dev + o_sj);
        a.n_jobs = nj; a.n_tiles = tiles; a.any_apply = 1;
        a.thresholds = reinterpret_cast<float*>(dev + o_thr);
        a.counts = reinterpret_cast<long long*>(dev + o_cnt);
        rc = wt_select_launch(a, stream);
        if (rc) break;
        const MwJob* dj = reinterpret_cast<const MwJob*>(dev + o_mj);
        unsigned int* teq = reinterpret_cast<unsigned int*>(dev + o_teq);
        unsigned long long* tpx = reinterpret_cast<unsigned long long*>(dev + o_tpx);
        unsigned long long* less = reinterpret_cast<unsigned long long*>(dev + o_less);
        unsigned long long* dnz = reinterpret_cast<unsigned long long*>(dev + o_nz);
        mw_count_kernel<<<tiles, MW_THREADS, 0, stream>>>(dj, nj, a.state, teq, less);
        mw_scan_kernel<<<nj, 256, 0, stream>>>(dj, a.state, teq, tpx);
        mw_apply_kernel<<<tiles, MW_THREADS, 0, stream>>>(dj, nj, a.state, tpx, less, dnz);
        mw_finalize_kernel<<<(nj + 127) / 128, 128, 0, stream>>>(dj, nj, a.state, a.counts, dnz);
        MW_TRY(cudaGetLastError());
        MW_TRY(cudaMemcpyAsync(nz.data(), dnz, nj * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
        MW_TRY(cudaStreamSynchronize(stream));
#undef MW_TRY
        for (int j = 0; j < nj; ++j) nonzero_out[owner[j]] = (int64_t)nz[j];
    } while (0);
    if (rc) cudaStreamSynchronize(stream);
    return rc;
}

extern "C" int wt_scatter_zero_f32(float* w, int64_t n, const int64_t* idx, int64_t k, int64_t* nonzero_out, void* stream_) {
    WT_CHECK(n >= 0 && k >= 0, WT_E_INVALID, "negative size");
    WT_CHECK(nonzero_out != nullptr, WT_E_INVALID, "nonzero_out is NULL");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    WT_CUDA(cudaMemsetAsync(nonzero_out, 0, sizeof(int64_t), stream));
    if (n == 0) return WT_OK;
    WT_CHECK(w != nullptr && (k == 0 || idx != nullptr), WT_E_INVALID, "NULL argument");
    if (k > 0) scatter_zero_kernel<<<grid_for(k, 256, 8), 256, 0, stream>>>(w, n, reinterpret_cast<const long long*>(idx), k);
    count_kernel<<<grid_for(n, 256, 8), 256, 0, stream>>>(w, n, 0.0f, reinterpret_cast<unsigned long long*>(nonzero_out), 0);
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

extern "C" int wt_count_f32(const float* x, int64_t n, float eps, int64_t* out, void* stream_) {
    WT_CHECK(out != nullptr && n >= 0, WT_E_INVALID, "bad argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    WT_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(int64_t), stream));
    if (n == 0) return WT_OK;
    WT_CHECK(x != nullptr, WT_E_INVALID, "x is NULL");
    count_kernel<<<grid_for(n, 256, 8), 256, 0, stream>>>(x, n, eps, reinterpret_cast<unsigned long long*>(out), 1);
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}
