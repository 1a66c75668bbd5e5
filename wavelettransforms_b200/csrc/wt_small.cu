// wt_small.cu -- small-plane regime of the DWT prune path (kH x kW conv-kernel planes).
//
// Replaces pywt.wavedec2 / pywt.waverec2 over axes (-2,-1) of a conv weight [O,I,kH,kW]
// (/root/reference/ResNet/dwt_pruning.py:67-68,77; dwt_pruning_NoEntropy.py:41,49) for the plane
// sizes conv layers have (1x1, 3x3, 5x5, 7x7), mode 'periodization', any of the 17 wavelets,
// levels 0..WT_SMALL_MAX_LEVEL -- including N < F where every tap wraps several times (F6).
//
// One thread owns one plane; every level of the decomposition lives in registers (all loop bounds
// are template constants, so after unrolling every array index is static and every filter tap is a
// constant-bank operand taken from the __grid_constant__ kernel parameter).  A CTA's planes are one
// contiguous span of global memory: it is staged through shared memory with a 1-D TMA bulk copy
// (cp.async.bulk + mbarrier) and read back at an odd word stride (9, 25, 49 floats: conflict free).
// Coefficients are stored coefficient-major ([ncoef][batch]) so warps read and write them coalesced.  (A per-CTA
// block-interleaved layout [batch / 128][ncoef][128] evens out the select's work per tile but costs more in these
// kernels' DRAM locality than it saves there: ResNet-152 1.19 -> 1.42 ms.  Measured and dropped.)
//
// Arithmetic follows oracle/dwt_oracle.c exactly (fp32, multiply then add, no FMA, same order).
//
// This file is compiled twice: as is (faithful arithmetic: multiply, then add, no FMA) and through wt_small_fma.cu with
// WT_SMALL_FMA = 1 (the named fast mode `fma_taporder`: the same tap order, every multiply-add fused; oracle twin:
// orc_set_arithmetic(1)).
#include "wt_common.cuh"
#include "wt_small_core.cuh"

#ifndef WT_SMALL_FMA
#define WT_SMALL_FMA 0
#endif

namespace {

using namespace wtsmall;
constexpr bool kFma = WT_SMALL_FMA != 0;

__device__ __forceinline__ int find_small_job(const WtSmallJob* __restrict__ jobs, int n_jobs, int tile) {
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (__ldg(&jobs[mid].tile_begin) <= tile) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// ---- TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP) -------------------------------------------
__device__ __forceinline__ unsigned int smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, unsigned int bytes, unsigned long long* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned int parity) {
    unsigned int ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, unsigned int bytes) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// stage `cnt` planes (contiguous) into shared memory
template <int HW>
__device__ __forceinline__ void stage_in(float* s, const float* __restrict__ g, int cnt, unsigned long long* bar) {
    const unsigned int bytes = (unsigned int)cnt * HW * 4u;
    const bool tma_ok = ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) && ((bytes & 15u) == 0);
    if (tma_ok) {
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            bulk_load(s, g, bytes, bar);
        }
        __syncthreads();
        mbar_wait(bar, 0);
    } else {
        for (int i = threadIdx.x; i < cnt * HW; i += WT_SMALL_THREADS) s[i] = __ldg(&g[i]);
        __syncthreads();
    }
}

template <int HW>
__device__ __forceinline__ void stage_out(float* g, const float* s, int cnt) {
    const unsigned int bytes = (unsigned int)cnt * HW * 4u;
    const bool tma_ok = ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) && ((bytes & 15u) == 0);
    __syncthreads();
    if (tma_ok) {
        if (threadIdx.x == 0) bulk_store(g, s, bytes);
    } else {
        for (int i = threadIdx.x; i < cnt * HW; i += WT_SMALL_THREADS) g[i] = s[i];
    }
}

// Stores the CTA's staged output tile at the same slot of every peer's buffer over NVLink.  One elected thread hands the
// copies to the TMA engine as bulk stores from the shared-memory tile the kernel holds anyway (cp.async.bulk, one per
// peer), so the SM does not sit on NVLink stores: it only waits until the engine has READ the tile (wait_group.read)
// before the CTA retires; the writes drain on their own and are complete when the kernel is.  (Round 1 issued plain
// 16-byte stores from all threads: 7 sequential peer stores per element stalled the LSU, the peer-store mode won only
// 0.10 ms over the all-gather at 8 GPUs.)  Misaligned tiles fall back to the thread stores.
__device__ __noinline__ void mirror_tile(const float* s, float* out, const WtPeers* __restrict__ peers, int nel) {
    const int n = __ldg(&peers->n);
    const unsigned int bytes = (unsigned int)nel * 4u;
    if (__ldg(&peers->mc)) {   // NVLS: one store per value to the multicast mapping, the switch replicates it to every rank
        char* g = reinterpret_cast<char*>(out) + __ldg(&peers->delta[0]);
        if (((reinterpret_cast<uintptr_t>(g) & 15u) == 0) && (nel & 3) == 0) {
            for (int i = threadIdx.x; i < (nel >> 2); i += WT_SMALL_THREADS)
                wt_mc_store4(g + (size_t)i * 16, reinterpret_cast<const float4*>(s)[i]);
        } else {
            for (int i = threadIdx.x; i < nel; i += WT_SMALL_THREADS) wt_mc_store1(g + (size_t)i * 4, s[i]);
        }
        return;
    }
    if (((reinterpret_cast<uintptr_t>(out) & 15u) == 0) && (bytes & 15u) == 0) {   // deltas are multiples of 16 (wt_plan_set_peers)
        if (threadIdx.x == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            for (int q = 0; q < n; ++q) {
                char* g = reinterpret_cast<char*>(out) + __ldg(&peers->delta[q]);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        return;
    }
    for (int q = 0; q < n; ++q) {
        float* g = reinterpret_cast<float*>(reinterpret_cast<char*>(out) + __ldg(&peers->delta[q]));
        for (int i = threadIdx.x; i < nel; i += WT_SMALL_THREADS) g[i] = s[i];
    }
}

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
// MODE 0: wavedec2 of every plane, coefficients to the coefficient-major workspace.
// MODE 1: the same, and -- for the precompacted select (wt_select.cu) -- every coefficient is held against the job's bracket
//         while it is in a register: keys below it are counted, keys inside it are appended to the select job's candidate
//         list (warp-private staging, one global atomic per ~224 keys), so the select never streams the coefficients.
// MODE 2: the sample pass that precedes it: one tile out of every 2^shift of the job, 12-bit histogram of |c| (key bits
//         30..19) into the select job's hist1 / hist1c; nothing else is written.
constexpr int SM_WFLUSH = 224, SM_WSTAGE = SM_WFLUSH + 32;

template <int H, int W, int F, class Z, int MODE>
__global__ void __launch_bounds__(WT_SMALL_THREADS)
wt_small_analysis_kernel(const __grid_constant__ WtBank bk, const WtSmallJob* __restrict__ jobs, int n_jobs,
                         long long* __restrict__ counts, WtSelState* state, unsigned long long* hist1, unsigned long long* hist1c) {
    constexpr int HW = H * W;
    __shared__ __align__(16) float s_in[WT_SMALL_THREADS * HW];
    __shared__ unsigned long long bar;
    __shared__ unsigned int s_cnt[WT_SMALL_THREADS / 32];
    __shared__ unsigned int s_x[MODE == 2 ? WT_SEL_BINS1 : 1];
    const int tile = blockIdx.x;
    const int j = find_small_job(jobs, n_jobs, tile);
    const WtSmallJob job = jobs[j];
    if (MODE == 2) {   // CTA-uniform early exits: jobs without a select, tiles outside the sample
        const int s = job.sample_shift;
        if (s < 0 || job.sel_job < 0) return;
        if (s > 0) {
            const int t = tile - job.tile_begin, g = t >> s;
            int avail = job.n_tiles - (g << s);
            if (avail > (1 << s)) avail = 1 << s;
            if (t != (g << s) + (int)(((unsigned int)g * 5u + 3u) % (unsigned int)avail)) return;
        }
        for (int b = threadIdx.x; b < WT_SEL_BINS1; b += WT_SMALL_THREADS) s_x[b] = 0u;
    }
    const long long p0 = (long long)(tile - job.tile_begin) * WT_SMALL_THREADS;
    const long long left = job.batch - p0;
    const int cnt = left < WT_SMALL_THREADS ? (int)left : WT_SMALL_THREADS;
    stage_in<HW>(s_in, job.in + p0 * HW, cnt, &bar);
    const bool valid = (int)threadIdx.x < cnt;
    unsigned int nz = 0;
    float x[HW];
#pragma unroll
    for (int i = 0; i < HW; ++i) {
        x[i] = valid ? s_in[threadIdx.x * HW + i] : 0.0f;   // lanes without a plane run the transform on zeros: the warp votes below need them
        nz += (x[i] != 0.0f);
    }
    if (MODE == 2) {
        auto put = [&](int, float v) {
            if (valid) atomicAdd(&s_x[(__float_as_uint(v) & 0x7fffffffu) >> 19], 1u);
        };
        SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::analyze(x, job.level, bk, put);
        __syncthreads();
        unsigned long long* gh = hist1 + (size_t)job.sel_job * WT_SEL_BINS1;
        unsigned long long* gc = hist1c + (size_t)job.sel_job * WT_SEL_COARSE;
        for (int b = threadIdx.x; b < WT_SEL_BINS1; b += WT_SMALL_THREADS) {   // a warp covers 32 consecutive fine bins = one coarse bin
            const unsigned int v = s_x[b];
            if (v) atomicAdd(&gh[b], (unsigned long long)v);
            unsigned int sum = v;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, d);
            if ((threadIdx.x & 31) == 0 && sum) atomicAdd(&gc[b >> 5], (unsigned long long)sum);
        }
        return;
    }
    float* __restrict__ cbase = job.coef + p0 + threadIdx.x;
    const long long stride = job.batch;
    {
        auto put = [&](int idx, float v) {
            if (valid) cbase[(long long)idx * stride] = v;
        };
        SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::analyze(x, job.level, bk, put);
    }
    // nonzeros of the input (dwt_pruning_NoEntropy.py:58) and numel
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) nz += __shfl_down_sync(0xffffffffu, nz, d);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = nz;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < WT_SMALL_THREADS / 32; ++w) t += s_cnt[w];
        atomicAdd(reinterpret_cast<unsigned long long*>(counts + (size_t)job.layer * 4 + 3), t);
        if (tile == job.tile_begin) counts[(size_t)job.layer * 4] = job.batch * HW;
    }
}

// MODE 0 with the next tile's copy in flight: a CTA takes SM_PAIR consecutive tiles, issues the TMA loads of all of them up
// front (one mbarrier each) and transforms them one after the other, so only the first load's latency is exposed (ncu,
// round 2: the one-tile kernel issues 75 % of the time; its four warps all wait for the tile before any of them computes).
#ifndef WT_SMALL_PAIR
#define WT_SMALL_PAIR 2
#endif
constexpr int SM_PAIR = WT_SMALL_PAIR;
// tiles per CTA for H x W planes: as many as fit 40 KB of static shared memory (7 x 7 planes: one)
// (1 x 1 planes -- 512 bytes per tile, nothing but memory traffic -- take eight: measured on ResNet-152's 1 x 1 class 330 us with
// one tile per CTA, 271 with two, 250 with four; the flop-bound 3 x 3 classes lose 3 us beyond two)
template <int H, int W> struct PairTiles {
    static constexpr int value = (H * W == 1) ? 8 : ((H * W * WT_SMALL_THREADS * 4 * SM_PAIR <= 40960) ? SM_PAIR : 1);
};

template <int H, int W, int F, class Z>
__global__ void __launch_bounds__(WT_SMALL_THREADS)
wt_small_analysis_pair_kernel(const __grid_constant__ WtBank bk, const WtSmallJob* __restrict__ jobs, int n_jobs, int n_tiles,
                              long long* __restrict__ counts) {
    constexpr int HW = H * W, NP = PairTiles<H, W>::value;
    __shared__ __align__(16) float s_in[NP][WT_SMALL_THREADS * HW];
    __shared__ unsigned long long bar[NP];
    __shared__ unsigned int s_cnt[WT_SMALL_THREADS / 32];
    __shared__ int s_job[NP], s_planes[NP], s_tma[NP];
    __shared__ long long s_p0[NP];
    const int tile0 = blockIdx.x * NP;
    if (threadIdx.x < NP) {   // NP <= 8 threads describe one tile each and start its copy
        const int k = threadIdx.x, tile = tile0 + k;
        s_planes[k] = 0;
        if (tile < n_tiles) {
            const int j = find_small_job(jobs, n_jobs, tile);
            const long long batch = jobs[j].batch, p0 = (long long)(tile - jobs[j].tile_begin) * WT_SMALL_THREADS;
            const long long left = batch - p0;
            const int cnt = left < WT_SMALL_THREADS ? (int)left : WT_SMALL_THREADS;
            const float* src = jobs[j].in + p0 * HW;
            const unsigned int bytes = (unsigned int)cnt * HW * 4u;
            const bool tma_ok = ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) && ((bytes & 15u) == 0);
            s_job[k] = j; s_planes[k] = cnt; s_p0[k] = p0; s_tma[k] = tma_ok ? 1 : 0;
            if (tma_ok) {
                mbar_init(&bar[k], 1);
                bulk_load(s_in[k], src, bytes, &bar[k]);
            }
        }
    }
    __syncthreads();
#pragma unroll 1
    for (int k = 0; k < NP; ++k) {
        const int cnt = s_planes[k];
        if (cnt == 0) break;   // CTA-uniform: past the last tile of the class
        const WtSmallJob job = jobs[s_job[k]];
        const long long p0 = s_p0[k];
        if (s_tma[k]) {
            mbar_wait(&bar[k], 0);
        } else {
            const float* __restrict__ g = job.in + p0 * HW;
            for (int i = threadIdx.x; i < cnt * HW; i += WT_SMALL_THREADS) s_in[k][i] = __ldg(&g[i]);
            __syncthreads();
        }
        const bool valid = (int)threadIdx.x < cnt;
        unsigned int nz = 0;
        float x[HW];
#pragma unroll
        for (int i = 0; i < HW; ++i) {
            x[i] = valid ? s_in[k][threadIdx.x * HW + i] : 0.0f;
            nz += (x[i] != 0.0f);
        }
        float* __restrict__ cbase = job.coef + p0 + threadIdx.x;
        const long long stride = job.batch;
        auto put = [&](int idx, float v) {
            if (valid) cbase[(long long)idx * stride] = v;
        };
        SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::analyze(x, job.level, bk, put);
        // nonzeros of the input (dwt_pruning_NoEntropy.py:58) and numel
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) nz += __shfl_down_sync(0xffffffffu, nz, d);
        if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = nz;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long t = 0;
            for (int w = 0; w < WT_SMALL_THREADS / 32; ++w) t += s_cnt[w];
            atomicAdd(reinterpret_cast<unsigned long long*>(counts + (size_t)job.layer * 4 + 3), t);
            if (p0 == 0) counts[(size_t)job.layer * 4] = job.batch * HW;
        }
        __syncthreads();   // s_cnt is reused by the next tile
    }
}

// MODE 1 as a kernel of its own: a CTA walks SM_TPC consecutive tiles of the class and carries its warp-private staging and
// its counters from tile to tile, so the job's global counters see one atomic per CTA and job instead of one per warp and
// tile (measured: with one tile per CTA the returning atomics on s_count made the haar analysis 3.7x slower).
constexpr int SM_TPC = 8;

template <int H, int W, int F, class Z>
__global__ void __launch_bounds__(WT_SMALL_THREADS)
wt_small_analysis_compact_kernel(const __grid_constant__ WtBank bk, const WtSmallJob* __restrict__ jobs, int n_jobs, int n_tiles,
                                 long long* __restrict__ counts, WtSelState* state) {
    constexpr int HW = H * W;
    __shared__ __align__(16) float s_in[WT_SMALL_THREADS * HW];
    __shared__ unsigned long long bar;
    __shared__ unsigned int s_cnt[WT_SMALL_THREADS / 32];
    __shared__ unsigned int s_x[(WT_SMALL_THREADS / 32) * SM_WSTAGE];
    const int tile0 = blockIdx.x * SM_TPC, tile1 = min(n_tiles, tile0 + SM_TPC);
    const int lane = threadIdx.x & 31;
    const unsigned int lt = (1u << lane) - 1u;
    unsigned int* wbuf = s_x + (threadIdx.x >> 5) * SM_WSTAGE;
    int j = find_small_job(jobs, n_jobs, tile0);
    WtSmallJob job = jobs[j];
    WtSelState* st = nullptr;
    unsigned int lo = 0, width = 0, wc = 0, nb = 0, nz = 0;   // wc is warp-uniform
    auto open_job = [&]() {
        st = job.sel_job >= 0 ? state + job.sel_job : nullptr;
        const unsigned int n_bins = st ? __ldcg(&st->n_bins) : 0u, lo_bin = st ? __ldcg(&st->lo_bin) : 0u;   // CTA-uniform
        // a bracket that starts at bin 0 leaves the exact zeros out of the list (wt_select.cu, phase_s1_compact)
        lo = (lo_bin << 19) | (lo_bin == 0u ? 1u : 0u);
        width = n_bins ? ((lo_bin + n_bins) << 19) - lo : 0u;
    };
    auto flush = [&]() {
        unsigned int gbase = 0;
        if (lane == 0) gbase = atomicAdd(&st->s_count, wc);
        gbase = __shfl_sync(0xffffffffu, gbase, 0);
        __syncwarp();
        for (unsigned int i = lane; i < wc; i += 32)
            if (gbase + i < job.cand_cap) job.cand[gbase + i] = wbuf[i];
        __syncwarp();
        wc = 0;
    };
    auto close_job = [&]() {   // the CTA's share of the job: staged keys, keys below the bracket, nonzero inputs
        if (width) {
            if (wc) flush();
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) nb += __shfl_down_sync(0xffffffffu, nb, d);
            if (lane == 0 && nb) atomicAdd(&st->below, (unsigned long long)nb);
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) nz += __shfl_down_sync(0xffffffffu, nz, d);
        if (lane == 0 && nz) atomicAdd(reinterpret_cast<unsigned long long*>(counts + (size_t)job.layer * 4 + 3), (unsigned long long)nz);
        nb = nz = 0;
    };
    open_job();
    for (int tile = tile0; tile < tile1; ++tile) {
        if (tile >= job.tile_begin + job.n_tiles) {
            close_job();
            job = jobs[++j];
            open_job();
        }
        __syncthreads();   // every thread has its plane of the previous tile in registers: s_in and the mbarrier are free
        const long long p0 = (long long)(tile - job.tile_begin) * WT_SMALL_THREADS;
        const long long left = job.batch - p0;
        const int cnt = left < WT_SMALL_THREADS ? (int)left : WT_SMALL_THREADS;
        stage_in<HW>(s_in, job.in + p0 * HW, cnt, &bar);
        const bool valid = (int)threadIdx.x < cnt;
        float x[HW];
#pragma unroll
        for (int i = 0; i < HW; ++i) {
            x[i] = valid ? s_in[threadIdx.x * HW + i] : 0.0f;   // lanes without a plane run the transform on zeros: the warp votes need them
            nz += (x[i] != 0.0f);
        }
        float* __restrict__ cbase = job.coef + p0 + threadIdx.x;
        const long long stride = job.batch;
        auto put = [&](int idx, float v) {
            if (valid) cbase[(long long)idx * stride] = v;
            if (width) {
                const unsigned int key = __float_as_uint(v) & 0x7fffffffu;
                nb += (valid && key < lo);
                const bool hit = valid && (key - lo < width);
                const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
                if (ballot) {
                    if (hit) wbuf[wc + __popc(ballot & lt)] = key;
                    wc += __popc(ballot);
                    if (wc > SM_WFLUSH) flush();
                }
            }
        };
        SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::analyze(x, job.level, bk, put);
        if (tile == job.tile_begin && threadIdx.x == 0) counts[(size_t)job.layer * 4] = job.batch * HW;
    }
    close_job();
}

template <int H, int W, int F, class Z>
__global__ void __launch_bounds__(WT_SMALL_THREADS)
wt_small_synthesis_kernel(const __grid_constant__ WtBank bk, const WtSmallJob* __restrict__ jobs, int n_jobs,
                          const float* __restrict__ thresholds, long long* __restrict__ counts) {
    constexpr int HW = H * W;
    __shared__ __align__(16) float s_out[WT_SMALL_THREADS * HW];
    __shared__ unsigned int s_cnt[WT_SMALL_THREADS / 32];
    const int tile = blockIdx.x;
    const int j = find_small_job(jobs, n_jobs, tile);
    const WtSmallJob job = jobs[j];
    const long long p0 = (long long)(tile - job.tile_begin) * WT_SMALL_THREADS;
    const long long left = job.batch - p0;
    const int cnt = left < WT_SMALL_THREADS ? (int)left : WT_SMALL_THREADS;
    const float t = job.use_abs ? job.abs_threshold : __ldg(&thresholds[job.layer]);
    unsigned int zc = 0;
    if ((int)threadIdx.x < cnt) {
        float x[HW];
        const float* __restrict__ cbase = job.coef + p0 + threadIdx.x;
        const long long stride = job.batch;
        auto get = [&](int idx) {
            const float c = __ldg(&cbase[(long long)idx * stride]);
            return (fabsf(c) < t) ? 0.0f : c;  // np.where(|c| < t, 0, c), strict (dwt_pruning.py:31)
        };
        SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::synthesize(x, job.level, bk, get);
#pragma unroll
        for (int i = 0; i < HW; ++i) {
            s_out[threadIdx.x * HW + i] = x[i];
            zc += (x[i] == 0.0f);
        }
    }
    stage_out<HW>(job.out + p0 * HW, s_out, cnt);
    // layer-sharded run: the staged tile also goes to the same slot of every peer's buffer (out of line and through the
    // job table in global memory, so the single-GPU kernel keeps its registers and instruction stream)
    if (__ldg(&jobs[j].peers.n)) mirror_tile(s_out, job.out + p0 * HW, &jobs[j].peers, cnt * HW);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) zc += __shfl_down_sync(0xffffffffu, zc, d);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = zc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long z = 0;
        for (int w = 0; w < WT_SMALL_THREADS / 32; ++w) z += s_cnt[w];
        atomicAdd(reinterpret_cast<unsigned long long*>(counts + (size_t)job.layer * 4 + 2), z);
        atomicAdd(reinterpret_cast<unsigned long long*>(counts + (size_t)job.layer * 4 + 1),
                  (unsigned long long)cnt * HW - z);
    }
}

// zero-tap masks of the actual float32 taps
struct TapMasks {
    unsigned dl, dh, rl, rh;
};
TapMasks tap_masks(const WtBank& b) {
    TapMasks m = {0, 0, 0, 0};
    for (int j = 0; j < b.F; ++j) {
        if (b.dec_lo[j] == 0.0f) m.dl |= 1u << j;
        if (b.dec_hi[j] == 0.0f) m.dh |= 1u << j;
        if (b.rec_lo[j] == 0.0f) m.rl |= 1u << j;
        if (b.rec_hi[j] == 0.0f) m.rh |= 1u << j;
    }
    return m;
}

// (F, masks) of the biorthogonal banks with zero-padded filters: bior1.3, bior2.2, bior4.4, rbio1.3, rbio2.2, rbio4.4
#define WT_ZERO_TAP_BANKS(X)          \
    X(6, 0x0, 0x33, 0x33, 0x0)        \
    X(6, 0x1, 0x31, 0x31, 0x1)        \
    X(10, 0x1, 0x301, 0x301, 0x1)     \
    X(6, 0x33, 0x0, 0x0, 0x33)        \
    X(6, 0x23, 0x20, 0x20, 0x23)      \
    X(10, 0x203, 0x200, 0x200, 0x203)

template <int H, int W, int MODE>
int launch_analysis_hw(const WtBank& bank, const WtSmallJob* jobs, int n_jobs, int n_tiles, long long* counts, WtSelState* state,
                       unsigned long long* hist1, unsigned long long* hist1c, cudaStream_t stream) {
    const TapMasks m = tap_masks(bank);
#define WT_SPEC(FL, DL, DH, RL, RH)                                                                                       \
    if (bank.F == FL && m.dl == DL && m.dh == DH && m.rl == RL && m.rh == RH) {                                         \
        if (MODE == 1)                                                                                                   \
            wt_small_analysis_compact_kernel<H, W, FL, ZeroTaps<DL, DH, RL, RH, kFma>>                                   \
                <<<(n_tiles + SM_TPC - 1) / SM_TPC, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, n_tiles, counts, state); \
        else if (MODE == 0)                                                                                              \
            wt_small_analysis_pair_kernel<H, W, FL, ZeroTaps<DL, DH, RL, RH, kFma>>                                      \
                <<<(n_tiles + PairTiles<H, W>::value - 1) / PairTiles<H, W>::value, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, n_tiles, counts); \
        else                                                                                                             \
            wt_small_analysis_kernel<H, W, FL, ZeroTaps<DL, DH, RL, RH, kFma>, 2>                                        \
                <<<n_tiles, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, counts, state, hist1, hist1c);             \
        WT_CUDA(cudaGetLastError());                                                                                     \
        return WT_OK;                                                                                                    \
    }
    WT_ZERO_TAP_BANKS(WT_SPEC)
#undef WT_SPEC
#define WT_CASE(FL)                                                                                                       \
    case FL:                                                                                                              \
        if (MODE == 1)                                                                                                    \
            wt_small_analysis_compact_kernel<H, W, FL, ZeroTaps<0, 0, 0, 0, kFma>>                                        \
                <<<(n_tiles + SM_TPC - 1) / SM_TPC, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, n_tiles, counts, state); \
        else if (MODE == 0)                                                                                               \
            wt_small_analysis_pair_kernel<H, W, FL, ZeroTaps<0, 0, 0, 0, kFma>>                                           \
                <<<(n_tiles + PairTiles<H, W>::value - 1) / PairTiles<H, W>::value, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, n_tiles, counts); \
        else                                                                                                              \
            wt_small_analysis_kernel<H, W, FL, ZeroTaps<0, 0, 0, 0, kFma>, 2>                                             \
                <<<n_tiles, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, counts, state, hist1, hist1c);              \
        break;
    switch (bank.F) {
        WT_CASE(2) WT_CASE(4) WT_CASE(6) WT_CASE(8) WT_CASE(10) WT_CASE(12) WT_CASE(18)
        default:
            wt_set_error("small-plane path: unsupported filter length %d", bank.F);
            return WT_E_UNSUPPORTED;
    }
#undef WT_CASE
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

template <int H, int W>
int launch_synthesis_hw(const WtBank& bank, const WtSmallJob* jobs, int n_jobs, int n_tiles, const float* thresholds,
                        long long* counts, cudaStream_t stream) {
    const TapMasks m = tap_masks(bank);
#define WT_SPEC(FL, DL, DH, RL, RH)                                                                                       \
    if (bank.F == FL && m.dl == DL && m.dh == DH && m.rl == RL && m.rh == RH) {                                         \
        wt_small_synthesis_kernel<H, W, FL, ZeroTaps<DL, DH, RL, RH, kFma>>                                                    \
            <<<n_tiles, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, thresholds, counts);                           \
        WT_CUDA(cudaGetLastError());                                                                                     \
        return WT_OK;                                                                                                    \
    }
    WT_ZERO_TAP_BANKS(WT_SPEC)
#undef WT_SPEC
#define WT_CASE(FL)                                                                                                       \
    case FL:                                                                                                              \
        wt_small_synthesis_kernel<H, W, FL, ZeroTaps<0, 0, 0, 0, kFma>>                                                                  \
            <<<n_tiles, WT_SMALL_THREADS, 0, stream>>>(bank, jobs, n_jobs, thresholds, counts);                           \
        break;
    switch (bank.F) {
        WT_CASE(2) WT_CASE(4) WT_CASE(6) WT_CASE(8) WT_CASE(10) WT_CASE(12) WT_CASE(18)
        default:
            wt_set_error("small-plane path: unsupported filter length %d", bank.F);
            return WT_E_UNSUPPORTED;
    }
#undef WT_CASE
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

}  // namespace

#if WT_SMALL_FMA
#define WT_SMALL_ENTRY(name) name##_fma
#else
#define WT_SMALL_ENTRY(name) name##_faithful
int wt_small_analysis_mode_launch_faithful(int H, int W, int mode, const WtBank& bank, const WtSmallJob* jobs_dev, int n_jobs,
                                           int n_tiles, long long* counts, WtSelState* state, unsigned long long* hist1,
                                           unsigned long long* hist1c, cudaStream_t stream);
int wt_small_analysis_mode_launch_fma(int H, int W, int mode, const WtBank& bank, const WtSmallJob* jobs_dev, int n_jobs,
                                      int n_tiles, long long* counts, WtSelState* state, unsigned long long* hist1,
                                      unsigned long long* hist1c, cudaStream_t stream);
int wt_small_synthesis_launch_faithful(int H, int W, const WtBank& bank, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                                       const float* thresholds, long long* counts, cudaStream_t stream);
int wt_small_synthesis_launch_fma(int H, int W, const WtBank& bank, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                                  const float* thresholds, long long* counts, cudaStream_t stream);

bool wt_small_supported(int H, int W, int F, int level, int mode) {
    if (mode != WT_MODE_PERIODIZATION || level < 0 || level > WT_SMALL_MAX_LEVEL) return false;
    if (H != W || !(H == 1 || H == 3 || H == 5 || H == 7)) return false;
    return F == 2 || F == 4 || F == 6 || F == 8 || F == 10 || F == 12 || F == 18;
}

int wt_small_ncoef(int H, int W, int level) {
    int n = 0, h = H, w = W;
    for (int l = 0; l < level; ++l) {
        h = (h + 1) / 2;
        w = (w + 1) / 2;
        n += 3 * h * w;
    }
    return n + h * w;
}

static int analysis_mode(int H, int W, int mode, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                         long long* counts, WtSelState* state, unsigned long long* hist1, unsigned long long* hist1c,
                         cudaStream_t stream) {
    if (fma) return wt_small_analysis_mode_launch_fma(H, W, mode, bank, jobs_dev, n_jobs, n_tiles, counts, state, hist1, hist1c, stream);
    return wt_small_analysis_mode_launch_faithful(H, W, mode, bank, jobs_dev, n_jobs, n_tiles, counts, state, hist1, hist1c, stream);
}
int wt_small_analysis_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                             long long* counts, cudaStream_t stream) {
    return analysis_mode(H, W, 0, bank, fma, jobs_dev, n_jobs, n_tiles, counts, nullptr, nullptr, nullptr, stream);
}
int wt_small_sample_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                           unsigned long long* hist1, unsigned long long* hist1c, cudaStream_t stream) {
    return analysis_mode(H, W, 2, bank, fma, jobs_dev, n_jobs, n_tiles, nullptr, nullptr, hist1, hist1c, stream);
}
int wt_small_analysis_compact_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                                     long long* counts, WtSelState* state, cudaStream_t stream) {
    return analysis_mode(H, W, 1, bank, fma, jobs_dev, n_jobs, n_tiles, counts, state, nullptr, nullptr, stream);
}
int wt_small_synthesis_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                              const float* thresholds, long long* counts, cudaStream_t stream) {
    if (fma) return wt_small_synthesis_launch_fma(H, W, bank, jobs_dev, n_jobs, n_tiles, thresholds, counts, stream);
    return wt_small_synthesis_launch_faithful(H, W, bank, jobs_dev, n_jobs, n_tiles, thresholds, counts, stream);
}
#endif

int WT_SMALL_ENTRY(wt_small_analysis_mode_launch)(int H, int W, int mode, const WtBank& bank, const WtSmallJob* jobs_dev, int n_jobs,
                                                  int n_tiles, long long* counts, WtSelState* state, unsigned long long* hist1,
                                                  unsigned long long* hist1c, cudaStream_t stream) {
    if (n_tiles <= 0) return WT_OK;
#define WT_HW(h, w)                                                                                                            \
    if (H == h && W == w) {                                                                                                    \
        if (mode == 0) return launch_analysis_hw<h, w, 0>(bank, jobs_dev, n_jobs, n_tiles, counts, state, hist1, hist1c, stream); \
        if (mode == 1) return launch_analysis_hw<h, w, 1>(bank, jobs_dev, n_jobs, n_tiles, counts, state, hist1, hist1c, stream); \
        return launch_analysis_hw<h, w, 2>(bank, jobs_dev, n_jobs, n_tiles, counts, state, hist1, hist1c, stream);             \
    }
    WT_HW(1, 1) WT_HW(3, 3) WT_HW(5, 5) WT_HW(7, 7)
#undef WT_HW
    wt_set_error("small-plane path: unsupported plane %dx%d", H, W);
    return WT_E_UNSUPPORTED;
}

int WT_SMALL_ENTRY(wt_small_synthesis_launch)(int H, int W, const WtBank& bank, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                                              const float* thresholds, long long* counts, cudaStream_t stream) {
    if (n_tiles <= 0) return WT_OK;
    if (H == 1 && W == 1) return launch_synthesis_hw<1, 1>(bank, jobs_dev, n_jobs, n_tiles, thresholds, counts, stream);
    if (H == 3 && W == 3) return launch_synthesis_hw<3, 3>(bank, jobs_dev, n_jobs, n_tiles, thresholds, counts, stream);
    if (H == 5 && W == 5) return launch_synthesis_hw<5, 5>(bank, jobs_dev, n_jobs, n_tiles, thresholds, counts, stream);
    if (H == 7 && W == 7) return launch_synthesis_hw<7, 7>(bank, jobs_dev, n_jobs, n_tiles, thresholds, counts, stream);
    wt_set_error("small-plane path: unsupported plane %dx%d", H, W);
    return WT_E_UNSUPPORTED;
}
