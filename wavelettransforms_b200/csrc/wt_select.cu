// wt_select.cu -- exact order-statistic select over |x| for a whole group of tensors in ONE
// cooperative launch, fused with the level-0 "select and zero" apply phase.
//
// Replaces np.percentile(np.abs(coeff_arr), p) + np.where(|c| < t, 0, c)
// (/root/reference/ResNet/dwt_pruning.py:27,31) and the zero count of dwt_pruning.py:88.
//
// Algorithm: 3-pass most-significant-digit radix select on the 31-bit pattern of |x|
// (12 + 10 + 9 bits).  The pattern order equals the numeric order for non-NaN floats.  Two ranks
// (k0 = floor(virtual index), k1 = k0 + 1) are tracked together so the interpolation of
// np.percentile can be reproduced from the exact sorted[k0] and sorted[k1].  Implicit zeros
// (coeffs_to_array padding) are added to digit 0 of every pass whose prefix is 0.
//
// Grid: persistent, one contiguous range of 4096-element tiles per CTA, all CTAs co-resident
// (cooperative launch); a grid barrier separates the passes.  Per pass each CTA histograms into
// shared memory, flushes the non-empty bins with global atomics, and the CTA that retires the last
// tile of a job scans that job's histogram (threadfence + ticket), so no extra barrier is needed.
// Pass 1 streams the tensors from HBM; passes 2, 3 and the apply phase re-read them from L2
// (126 MB on B200) when the group fits.
#include <cooperative_groups.h>

#include "wt_common.cuh"

namespace cg = cooperative_groups;

namespace {

__device__ __forceinline__ unsigned int abs_key(float v) { return __float_as_uint(v) & 0x7fffffffu; }

__device__ __forceinline__ int find_job(const WtSelJob* __restrict__ jobs, int n_jobs, int tile) {
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (__ldg(&jobs[mid].tile_begin) <= tile) lo = mid; else hi = mid - 1;
    }
    return lo;
}

template <int PASS>
__device__ __forceinline__ void bin_one(unsigned int key, unsigned int* sh, unsigned int p0, unsigned int p1, bool two) {
    if (PASS == 1) {
        atomicAdd(&sh[key >> 19], 1u);
    } else if (PASS == 2) {
        unsigned int top = key >> 19;
        if (top == p0) atomicAdd(&sh[(key >> 9) & 1023u], 1u);
        else if (two && top == p1) atomicAdd(&sh[1024u + ((key >> 9) & 1023u)], 1u);
    } else {
        unsigned int top = key >> 9;
        if (top == p0) atomicAdd(&sh[key & 511u], 1u);
        else if (two && top == p1) atomicAdd(&sh[512u + (key & 511u)], 1u);
    }
}

// exclusive block scan of one value per thread (512 threads), returns the exclusive prefix and the total
__device__ __forceinline__ unsigned long long block_excl_scan(unsigned long long v, unsigned long long* warp_sums,
                                                              unsigned long long* total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = (lane < WT_SEL_THREADS / 32) ? warp_sums[lane] : 0ull;
        unsigned long long winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        if (lane < WT_SEL_THREADS / 32) warp_sums[lane] = winc - w;  // exclusive warp offsets
        if (lane == WT_SEL_THREADS / 32 - 1) warp_sums[WT_SEL_THREADS / 32] = winc;
    }
    __syncthreads();
    unsigned long long excl = warp_sums[wid] + inc - v;
    *total = warp_sums[WT_SEL_THREADS / 32];
    __syncthreads();
    return excl;
}

// find the digit holding `rank` in hist[0..NB) (+ zeros in bin 0); result through shared memory
template <int NB>
__device__ void scan_find(const unsigned long long* hist, unsigned long long zeros, unsigned long long rank,
                          unsigned long long* warp_sums, unsigned int* res_digit, unsigned long long* res_rem) {
    constexpr int PER = (NB + WT_SEL_THREADS - 1) / WT_SEL_THREADS;
    unsigned long long c[PER];
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        int b = threadIdx.x * PER + i;
        unsigned long long v = (b < NB) ? __ldcg(&hist[b]) : 0ull;
        if (b == 0) v += zeros;
        c[i] = v;
        s += v;
    }
    unsigned long long total;
    unsigned long long excl = block_excl_scan(s, warp_sums, &total);
    if (rank >= total) rank = total ? total - 1 : 0;  // defensive; cannot happen for a consistent plan
    if (s > 0 && excl <= rank && rank < excl + s) {
        unsigned long long run = excl;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            if (rank < run + c[i]) {
                *res_digit = threadIdx.x * PER + i;
                *res_rem = rank - run;
                break;
            }
            run += c[i];
        }
    }
    __syncthreads();
}

__device__ __forceinline__ float lerp_threshold(float a, float b, double gamma, int mode) {
    if (mode == WT_LERP_LEGACY) {
        // numpy 1.22 _lerp on 0-d operands: (b - a) in float32, the rest in float64, then the
        // float32 array comparison demotes the float64 scalar.  No FMA.
        float d32 = __fsub_rn(b, a);
        double d = (double)d32, t;
        if (gamma >= 0.5) t = __dsub_rn((double)b, __dmul_rn(d, __dsub_rn(1.0, gamma)));
        else t = __dadd_rn((double)a, __dmul_rn(d, gamma));
        return __double2float_rn(t);
    } else {
        float g = (float)gamma;
        float d = __fsub_rn(b, a);
        if (g >= 0.5f) return __fsub_rn(b, __fmul_rn(d, __fsub_rn(1.0f, g)));
        return __fadd_rn(a, __fmul_rn(d, g));
    }
}

struct Shared {
    unsigned int hist[WT_SEL_BINS1];
    unsigned long long warp_sums[WT_SEL_THREADS / 32 + 1];
    unsigned long long res_rem[2];
    unsigned int res_digit[2];
    int is_last;
};

template <int PASS>
__device__ void flush_and_maybe_scan(const WtSelArgs& a, Shared& sh, int j, const WtSelJob& job, int my_tiles,
                                     unsigned int p0, unsigned int p1, bool two) {
    constexpr int NB = PASS == 1 ? WT_SEL_BINS1 : (PASS == 2 ? WT_SEL_BINS2 : WT_SEL_BINS3);
    constexpr int NBT = PASS == 1 ? NB : 2 * NB;
    unsigned long long* gh = PASS == 1 ? a.hist1 + (size_t)j * NB
                                       : (PASS == 2 ? a.hist2 + (size_t)j * 2 * NB : a.hist3 + (size_t)j * 2 * NB);
    __syncthreads();
    for (int b = threadIdx.x; b < NBT; b += WT_SEL_THREADS) {
        unsigned int v = sh.hist[b];
        if (v) atomicAdd(&gh[b], (unsigned long long)v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int ticket = atomicAdd(&a.done[j * 4 + (PASS - 1)], (unsigned int)my_tiles);
        sh.is_last = (ticket + (unsigned int)my_tiles == (unsigned int)job.n_tiles);
    }
    __syncthreads();
    if (!sh.is_last) return;
    __threadfence();
    // this CTA retired the job's last tile of the pass: locate the digit of both ranks
    WtSelState* st = &a.state[j];
    unsigned long long rank0 = PASS == 1 ? (unsigned long long)job.k0 : __ldcg(&st->rem[0]);
    unsigned long long rank1 = PASS == 1 ? (unsigned long long)job.k1 : __ldcg(&st->rem[1]);
    unsigned long long z0 = (PASS == 1 || p0 == 0u) ? (unsigned long long)job.zeros : 0ull;
    unsigned long long z1 = (PASS == 1 || p1 == 0u) ? (unsigned long long)job.zeros : 0ull;
    scan_find<NB>(gh, z0, rank0, sh.warp_sums, &sh.res_digit[0], &sh.res_rem[0]);
    scan_find<NB>(two ? gh + NB : gh, z1, rank1, sh.warp_sums, &sh.res_digit[1], &sh.res_rem[1]);
    if (threadIdx.x == 0) {
        constexpr int BITS = PASS == 1 ? 12 : (PASS == 2 ? 10 : 9);
        unsigned int n0 = PASS == 1 ? sh.res_digit[0] : ((p0 << BITS) | sh.res_digit[0]);
        unsigned int n1 = PASS == 1 ? sh.res_digit[1] : ((p1 << BITS) | sh.res_digit[1]);
        st->prefix[0] = n0;
        st->prefix[1] = n1;
        st->rem[0] = sh.res_rem[0];
        st->rem[1] = sh.res_rem[1];
        if (PASS == 3) {
            float av = __uint_as_float(n0), bv = __uint_as_float(n1);
            float t = lerp_threshold(av, bv, job.gamma, job.lerp_mode);
            st->a = av;
            st->b = bv;
            st->t = t;
            a.thresholds[job.layer] = t;
        }
        __threadfence();
    }
    __syncthreads();
}

template <int PASS>
__device__ void hist_phase(const WtSelArgs& a, Shared& sh, int t_begin, int t_end) {
    int cur = -1, my_tiles = 0;
    WtSelJob job;
    unsigned int p0 = 0, p1 = 0;
    bool two = false, vec = false;
    for (int tile = t_begin; tile < t_end; ++tile) {
        int j = find_job(a.jobs, a.n_jobs, tile);
        if (j != cur) {
            if (cur >= 0) flush_and_maybe_scan<PASS>(a, sh, cur, job, my_tiles, p0, p1, two);
            cur = j;
            my_tiles = 0;
            job = a.jobs[j];
            if (PASS > 1) {
                p0 = __ldcg(&a.state[j].prefix[0]);
                p1 = __ldcg(&a.state[j].prefix[1]);
                two = (p0 != p1);
            }
            vec = ((reinterpret_cast<uintptr_t>(job.x) & 15u) == 0);
            for (int b = threadIdx.x; b < WT_SEL_BINS1; b += WT_SEL_THREADS) sh.hist[b] = 0u;
            __syncthreads();
        }
        ++my_tiles;
        long long base = (long long)(tile - job.tile_begin) * WT_SEL_TILE;
        long long rem = job.n - base;
        const float* __restrict__ x = job.x + base;
        if (vec && rem >= WT_SEL_TILE) {
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float4 v0 = __ldg(&x4[threadIdx.x]);
            float4 v1 = __ldg(&x4[threadIdx.x + WT_SEL_THREADS]);
            bin_one<PASS>(abs_key(v0.x), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v0.y), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v0.z), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v0.w), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v1.x), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v1.y), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v1.z), sh.hist, p0, p1, two);
            bin_one<PASS>(abs_key(v1.w), sh.hist, p0, p1, two);
        } else {
            int cnt = rem < WT_SEL_TILE ? (int)rem : WT_SEL_TILE;
            for (int i = threadIdx.x; i < cnt; i += WT_SEL_THREADS)
                bin_one<PASS>(abs_key(__ldg(&x[i])), sh.hist, p0, p1, two);
        }
    }
    if (cur >= 0) flush_and_maybe_scan<PASS>(a, sh, cur, job, my_tiles, p0, p1, two);
}

__device__ __forceinline__ void block_add3(long long* dst, unsigned int zc, unsigned int nzb, unsigned int cnt,
                                           unsigned int* scratch) {
    // scratch: >= 3 * 16 unsigned ints of shared memory
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        zc += __shfl_down_sync(0xffffffffu, zc, d);
        nzb += __shfl_down_sync(0xffffffffu, nzb, d);
        cnt += __shfl_down_sync(0xffffffffu, cnt, d);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) { scratch[wid] = zc; scratch[16 + wid] = nzb; scratch[32 + wid] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long z = 0, nb = 0, c = 0;
        for (int w = 0; w < WT_SEL_THREADS / 32; ++w) { z += scratch[w]; nb += scratch[16 + w]; c += scratch[32 + w]; }
        atomicAdd(reinterpret_cast<unsigned long long*>(dst + 1), c - z);
        atomicAdd(reinterpret_cast<unsigned long long*>(dst + 2), z);
        atomicAdd(reinterpret_cast<unsigned long long*>(dst + 3), nb);
    }
    __syncthreads();
}

__device__ __forceinline__ float apply_one(float v, float t, unsigned int& zc, unsigned int& nzb) {
    float o = (fabsf(v) < t) ? 0.0f : v;
    zc += (o == 0.0f);
    nzb += (v != 0.0f);
    return o;
}

__device__ void apply_phase(const WtSelArgs& a, Shared& sh, int t_begin, int t_end) {
    int cur = -1;
    WtSelJob job;
    float t = 0.f;
    bool vec = false;
    unsigned int zc = 0, nzb = 0, cnt = 0;
    for (int tile = t_begin; tile < t_end; ++tile) {
        int j = find_job(a.jobs, a.n_jobs, tile);
        if (j != cur) {
            if (cur >= 0 && job.out) block_add3(a.counts + (size_t)job.layer * 4, zc, nzb, cnt, sh.hist);
            zc = nzb = cnt = 0;
            cur = j;
            job = a.jobs[j];
            t = __ldcg(&a.state[j].t);
            vec = (((reinterpret_cast<uintptr_t>(job.x) | reinterpret_cast<uintptr_t>(job.out)) & 15u) == 0);
        }
        if (!job.out) continue;
        if (tile == job.tile_begin && threadIdx.x == 0) a.counts[(size_t)job.layer * 4] = job.n;
        long long base = (long long)(tile - job.tile_begin) * WT_SEL_TILE;
        long long rem = job.n - base;
        const float* __restrict__ x = job.x + base;
        float* __restrict__ o = job.out + base;
        if (vec && rem >= WT_SEL_TILE) {
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float4* o4 = reinterpret_cast<float4*>(o);
            float4 v0 = x4[threadIdx.x];
            float4 v1 = x4[threadIdx.x + WT_SEL_THREADS];
            v0.x = apply_one(v0.x, t, zc, nzb); v0.y = apply_one(v0.y, t, zc, nzb);
            v0.z = apply_one(v0.z, t, zc, nzb); v0.w = apply_one(v0.w, t, zc, nzb);
            v1.x = apply_one(v1.x, t, zc, nzb); v1.y = apply_one(v1.y, t, zc, nzb);
            v1.z = apply_one(v1.z, t, zc, nzb); v1.w = apply_one(v1.w, t, zc, nzb);
            o4[threadIdx.x] = v0;
            o4[threadIdx.x + WT_SEL_THREADS] = v1;
            cnt += 8;
        } else {
            int n = rem < WT_SEL_TILE ? (int)rem : WT_SEL_TILE;
            for (int i = threadIdx.x; i < n; i += WT_SEL_THREADS) {
                o[i] = apply_one(x[i], t, zc, nzb);
                ++cnt;
            }
        }
    }
    if (cur >= 0 && job.out) block_add3(a.counts + (size_t)job.layer * 4, zc, nzb, cnt, sh.hist);
}

__global__ void __launch_bounds__(WT_SEL_THREADS) wt_select_kernel(WtSelArgs a, int do_apply) {
    __shared__ Shared sh;
    cg::grid_group grid = cg::this_grid();
    const long long T = a.n_tiles;
    const int t_begin = (int)((T * blockIdx.x) / gridDim.x);
    const int t_end = (int)((T * (blockIdx.x + 1)) / gridDim.x);
    hist_phase<1>(a, sh, t_begin, t_end);
    grid.sync();
    hist_phase<2>(a, sh, t_begin, t_end);
    grid.sync();
    hist_phase<3>(a, sh, t_begin, t_end);
    if (do_apply) {
        grid.sync();
        apply_phase(a, sh, t_begin, t_end);
    }
}

// out = |x| < t ? 0 : x with t from device (thr_dev) or host; counts4 += {., nonzero, zeros, nonzero_before}
__global__ void __launch_bounds__(256) wt_apply_threshold_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                                 long long n, const float* thr_dev, float thr_host,
                                                                 long long* counts4) {
    __shared__ unsigned int scratch[48];
    const float t = thr_dev ? __ldg(thr_dev) : thr_host;
    unsigned int zc = 0, nzb = 0, cnt = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        out[i] = apply_one(x[i], t, zc, nzb);
        ++cnt;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        zc += __shfl_down_sync(0xffffffffu, zc, d);
        nzb += __shfl_down_sync(0xffffffffu, nzb, d);
        cnt += __shfl_down_sync(0xffffffffu, cnt, d);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { scratch[wid] = zc; scratch[16 + wid] = nzb; scratch[32 + wid] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long z = 0, nb = 0, c = 0;
        for (int w = 0; w < 8; ++w) { z += scratch[w]; nb += scratch[16 + w]; c += scratch[32 + w]; }
        if (blockIdx.x == 0) counts4[0] = n;
        atomicAdd(reinterpret_cast<unsigned long long*>(counts4 + 1), c - z);
        atomicAdd(reinterpret_cast<unsigned long long*>(counts4 + 2), z);
        atomicAdd(reinterpret_cast<unsigned long long*>(counts4 + 3), nb);
    }
}

}  // namespace

size_t wt_select_scratch_bytes(int n_jobs) {
    size_t per = (size_t)WT_SEL_BINS1 * 8 + 2 * (size_t)WT_SEL_BINS2 * 8 + 2 * (size_t)WT_SEL_BINS3 * 8 + 4 * 4 +
                 sizeof(WtSelState);
    return per * (size_t)n_jobs + 256;
}

void wt_select_bind(WtSelArgs* a, void* scratch, int n_jobs) {
    char* p = static_cast<char*>(scratch);
    a->hist1 = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * WT_SEL_BINS1 * 8;
    a->hist2 = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * 2 * WT_SEL_BINS2 * 8;
    a->hist3 = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * 2 * WT_SEL_BINS3 * 8;
    a->state = reinterpret_cast<WtSelState*>(p); p += (size_t)n_jobs * sizeof(WtSelState);
    a->done = reinterpret_cast<unsigned int*>(p);
    a->n_jobs = n_jobs;
}

int wt_select_launch(const WtSelArgs& args, cudaStream_t stream) {
    static int max_blocks_per_sm = 0;
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    WT_CHECK(di.coop, WT_E_UNSUPPORTED, "device does not support cooperative launch");
    if (max_blocks_per_sm == 0) {
        int nb = 0;
        WT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wt_select_kernel, WT_SEL_THREADS, 0));
        WT_CHECK(nb > 0, WT_E_CUDA, "select kernel does not fit on an SM");
        max_blocks_per_sm = nb;
    }
    int grid = max_blocks_per_sm * di.sm_count;
    if (grid > args.n_tiles) grid = args.n_tiles;
    if (grid < 1) grid = 1;
    WtSelArgs a = args;
    int do_apply = a.any_apply;  // jobs without `out` skip their tiles inside the phase
    void* params[] = {&a, &do_apply};
    WT_CUDA(cudaLaunchCooperativeKernel((const void*)wt_select_kernel, dim3(grid), dim3(WT_SEL_THREADS), params, 0, stream));
    return WT_OK;
}

int wt_apply_threshold_launch(const float* x, float* out, long long n, const float* thr_dev, float thr_host,
                              long long* counts4, cudaStream_t stream) {
    if (n <= 0) return WT_OK;
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    long long blocks = (n + 255) / 256;
    long long cap = (long long)di.sm_count * 8;
    if (blocks > cap) blocks = cap;
    wt_apply_threshold_kernel<<<(int)blocks, 256, 0, stream>>>(x, out, n, thr_dev, thr_host, counts4);
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}
