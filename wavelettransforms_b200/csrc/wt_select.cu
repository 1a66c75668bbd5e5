// wt_select.cu -- exact order-statistic select over |x| for a whole group of tensors in ONE
// cooperative launch, fused with the level-0 "select and zero" apply phase.
//
// Replaces np.percentile(np.abs(coeff_arr), p) + np.where(|c| < t, 0, c)
// (/root/reference/ResNet/dwt_pruning.py:27,31) and the zero count of dwt_pruning.py:88.
//
// Keys are the 31-bit patterns of |x| (pattern order = numeric order for non-NaN floats).  Two
// ranks (k0 = floor(virtual index), k1 = k0 + 1) are tracked together so np.percentile's
// interpolation can be reproduced from the exact sorted[k0] and sorted[k1].  Implicit zeros
// (coeffs_to_array padding) are added to digit 0 wherever the prefix is 0.
//
// Persistent grid: every CTA owns one contiguous range of 4096-element tiles (cut into per-job segments), grid
// barriers separate the phases, and NO phase has a serial tail: after a barrier every CTA scans the (small,
// L2-resident) global histogram of its own job redundantly instead of waiting for one elected CTA to do it.
//
// SAMPLED PATH (launches of 40 Mi keys and more, and smaller ones whose candidate lists are predicted short, see
// wt_select_path / wt_select_predict_short): ONE full pass over the tensors, and that pass has no shared-memory atomics.
//   S0  12-bit histogram (key bits 30..19) of a SAMPLE: 4096 >> sample_shift elements of every tile, at a window
//       offset that rotates with the tile index; tensors below 32 Ki keys are read whole (the sample is exact)
//   S1  every CTA scans the sample histogram -> a bracket of consecutive 12-bit bins that holds both ranks
//       unless the sample lied (5 sigma of the binomial rank error on each side; exact for whole-tensor samples);
//       then the one streaming pass: count the keys below the bracket, compact the keys inside it into the job's
//       candidate list (warp-private staging, no CTA barrier, no histogram)
//   T   ONE CTA per job: verify (below <= rank < below + candidates, no overflow -- otherwise the job is handed to
//       the radix path); lists up to tail_max keys are finished right here by an MSD radix select with 11-bit
//       digits over the L2-resident list, CTA-local barriers only.  Longer lists raise big_flag and take the
//       grid-wide passes (a grid barrier, a flush and 296 redundant scans each):
//   C1  histogram of the top 12 bits of key - lo_key inside the bracket over the candidate list
//   C2  scan -> digit of both ranks; histogram of the next (7..11) bits of the matching candidates
//   C3  only when some job's bracket is wider than 16 bins (a rank next to a cluster of ties at zero: the bins in
//       between are many and nearly empty): the last <= 8 bits
//   P4  scan -> sorted[k0], sorted[k1] -> threshold; level-0 jobs: out = |x| < t ? 0 : x + counts
// The result is EXACT whatever the sample says: the bracket is only a guess of where to look, the counts decide.
//
// RADIX PATH (smaller launches, and any job the sampled path gave up on -- heavy ties away from zero --): MSD radix select, 12 + 10 + 9 key bits, three more phases before P4:
//   P1  histogram of key bits 30..19 (streams the tensors from HBM) -> global hist1
//   P2  scan hist1 -> digit of both ranks; histogram of bits 18..9 of the matching keys -> hist2,
//       and the matching keys (about 2 % of a Gaussian tensor) are compacted into the job's
//       candidate list through a shared-memory staging buffer (tensors re-read from L2)
//   P3  scan hist2 -> 22-bit prefixes; histogram of bits 8..0 over the CANDIDATE LIST only -> hist3
//       (a job whose candidates do not fit its list re-reads the tensor instead)
#include <cooperative_groups.h>

#include <cmath>
#include <cstdlib>
#include <cstring>

#include "wt_common.cuh"

namespace cg = cooperative_groups;

// longest candidate list the one-CTA tail takes when some other list needs the grid passes anyway (see wt_select_kernel;
// WT_SEL_TAIL_MAX in wt_common.cuh is the limit when none does).  Sweep 2048 .. 32768 on the headline: 8192 is the minimum.
#ifndef WT_SEL_TAIL_SHORT
#define WT_SEL_TAIL_SHORT 8192u
#endif

namespace {

constexpr int T = WT_SEL_THREADS;
constexpr int UNROLL = 4;  // float4 loads in flight per thread

__device__ __forceinline__ unsigned int abs_key(float v) { return __float_as_uint(v) & 0x7fffffffu; }

// tuning aid (WT_SELECT_STAMPS): 16 u64 slots per CTA; slots 0..5 phase boundaries, 6.. accumulated ns
__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
struct Acc {
    unsigned long long* slot;  // nullptr when stamps are off
    unsigned long long t0;
    __device__ __forceinline__ void begin() { if (slot && threadIdx.x == 0) t0 = now_ns(); }
    __device__ __forceinline__ void end(int i) { if (slot && threadIdx.x == 0) slot[i] += now_ns() - t0; }
    __device__ __forceinline__ void stamp(int i) { if (slot && threadIdx.x == 0) slot[i] = now_ns(); }
};

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

// The streaming loops are load -> wait -> process with every warp of the SM in step, so the memory system idles
// while the SM computes and vice versa; asking L2 for the next iteration's lines before processing this one
// overlaps the two at no register cost (a software-pipelined register prefetch spills under the 64-register cap).
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#ifndef WT_SEL_PREFETCH
#define WT_SEL_PREFETCH 1
#endif

// Calls op(float) on x[0..len): float4 loads, UNROLL of them in flight per thread.
template <class Op>
__device__ __forceinline__ void span_for_each(const float* x, long long len, Op op) {
    if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
        const long long nvec = len >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        for (long long v = threadIdx.x; v < nvec; v += (long long)T * UNROLL) {
            float4 r[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
                if (v + (long long)u * T < nvec) r[u] = __ldcg(&x4[v + (long long)u * T]);
            if (WT_SEL_PREFETCH) {
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if (v + (long long)(UNROLL + u) * T < nvec) prefetch_l2(&x4[v + (long long)(UNROLL + u) * T]);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
                if (v + (long long)u * T < nvec) { op(r[u].x); op(r[u].y); op(r[u].z); op(r[u].w); }
        }
        for (long long i = (nvec << 2) + threadIdx.x; i < len; i += T) op(__ldcg(&x[i]));
    } else {
        for (long long i = threadIdx.x; i < len; i += T) op(__ldcg(&x[i]));
    }
}

__device__ __forceinline__ unsigned long long load_bin(const unsigned long long* h, int b) { return __ldcg(&h[b]); }
__device__ __forceinline__ unsigned long long load_bin(const unsigned int* h, int b) { return h[b]; }  // CTA-private (shared)


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

constexpr unsigned int STAGE_FLUSH = 2048;                        // flush the staging buffer once it holds more keys
constexpr unsigned int STAGE = STAGE_FLUSH + T * UNROLL * 4;      // one more iteration always fits

struct Shared {
    unsigned int hist[WT_SEL_BINS1];
    unsigned int stage[STAGE];  // P2: candidate keys on their way to the job's global list, one slice per warp
    unsigned long long warp_sums[2][T / 32 + 1];
    unsigned long long res_rem[2];
    unsigned long long res_cnt[2];
    unsigned int res_digit[2];
};


struct Found {
    unsigned int d0, d1;          // digit holding rank0 / rank1
    unsigned long long r0, r1;    // rank inside that digit
    unsigned long long c0, c1;    // EXPLICIT keys in that digit (implicit zeros excluded)
};

// Locates both ranks in one go: h0 (+ z0 implicit zeros in bin 0) holds rank0, h1 (+ z1) holds rank1;
// h1 == h0 when the two ranks share a histogram.  Every thread returns the same result.
template <int NB, bool ALWAYS_SAME, class H>
__device__ Found scan_find_pair(const H* h0, const H* h1, unsigned long long z0, unsigned long long z1,
                                unsigned long long rank0, unsigned long long rank1, Shared& sh) {
    constexpr int PER = (NB + T - 1) / T;
    const bool same = ALWAYS_SAME || (h0 == h1);
    unsigned long long c0[PER], c1[ALWAYS_SAME ? 1 : PER];  // ALWAYS_SAME: c1 is never materialised
    unsigned long long s0 = 0, s1 = 0;
#define WT_C1(i) (ALWAYS_SAME ? c0[i] : c1[ALWAYS_SAME ? 0 : (i)])
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int b = threadIdx.x * PER + i;
        c0[i] = (b < NB) ? load_bin(h0, b) : 0ull;
        if (!ALWAYS_SAME) c1[ALWAYS_SAME ? 0 : i] = same ? 0ull : ((b < NB) ? load_bin(h1, b) : 0ull);
    }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        if (!ALWAYS_SAME && same) c1[ALWAYS_SAME ? 0 : i] = c0[i];
        if (threadIdx.x * PER + i == 0) {  // ALWAYS_SAME implies z0 == z1
            c0[i] += z0;
            if (!ALWAYS_SAME) c1[ALWAYS_SAME ? 0 : i] += z1;
        }
        s0 += c0[i];
        s1 += WT_C1(i);
    }
    // fused exclusive block scan of (s0, s1)
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long i0 = s0, i1 = s1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned long long o0 = __shfl_up_sync(0xffffffffu, i0, d), o1 = __shfl_up_sync(0xffffffffu, i1, d);
        if (lane >= d) { i0 += o0; i1 += o1; }
    }
    if (lane == 31) { sh.warp_sums[0][wid] = i0; sh.warp_sums[1][wid] = i1; }
    __syncthreads();
    if (wid == 0) {
        unsigned long long w0 = (lane < T / 32) ? sh.warp_sums[0][lane] : 0ull, w1 = (lane < T / 32) ? sh.warp_sums[1][lane] : 0ull;
        unsigned long long a0 = w0, a1 = w1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o0 = __shfl_up_sync(0xffffffffu, a0, d), o1 = __shfl_up_sync(0xffffffffu, a1, d);
            if (lane >= d) { a0 += o0; a1 += o1; }
        }
        if (lane < T / 32) { sh.warp_sums[0][lane] = a0 - w0; sh.warp_sums[1][lane] = a1 - w1; }
        if (lane == T / 32 - 1) { sh.warp_sums[0][T / 32] = a0; sh.warp_sums[1][T / 32] = a1; }
    }
    __syncthreads();
    const unsigned long long e0 = sh.warp_sums[0][wid] + i0 - s0, e1 = sh.warp_sums[1][wid] + i1 - s1;
    const unsigned long long tot0 = sh.warp_sums[0][T / 32], tot1 = sh.warp_sums[1][T / 32];
    if (rank0 >= tot0) rank0 = tot0 ? tot0 - 1 : 0;  // defensive; cannot happen for a consistent plan
    if (rank1 >= tot1) rank1 = tot1 ? tot1 - 1 : 0;
    if (s0 > 0 && e0 <= rank0 && rank0 < e0 + s0) {
        unsigned long long run = e0;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            if (rank0 < run + c0[i]) {
                const int b = threadIdx.x * PER + i;
                sh.res_digit[0] = b;
                sh.res_rem[0] = rank0 - run;
                sh.res_cnt[0] = c0[i] - (b == 0 ? z0 : 0ull);
                break;
            }
            run += c0[i];
        }
    }
    if (s1 > 0 && e1 <= rank1 && rank1 < e1 + s1) {
        unsigned long long run = e1;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            if (rank1 < run + WT_C1(i)) {
                const int b = threadIdx.x * PER + i;
                sh.res_digit[1] = b;
                sh.res_rem[1] = rank1 - run;
                sh.res_cnt[1] = WT_C1(i) - (b == 0 ? z1 : 0ull);
                break;
            }
            run += WT_C1(i);
        }
    }
    __syncthreads();
    Found f;
    f.d0 = sh.res_digit[0]; f.d1 = sh.res_digit[1];
    f.r0 = sh.res_rem[0]; f.r1 = sh.res_rem[1];
    f.c0 = sh.res_cnt[0]; f.c1 = sh.res_cnt[1];
    __syncthreads();  // the next call may overwrite the result slots
    return f;
#undef WT_C1
}

template <int PASS>
__device__ __forceinline__ unsigned long long* global_hist(const WtSelArgs& a, int j) {
    constexpr int NB = PASS == 1 ? WT_SEL_BINS1 : (PASS == 2 ? WT_SEL_BINS2 : WT_SEL_BINS3);
    return PASS == 1 ? a.hist1 + (size_t)j * NB : (PASS == 2 ? a.hist2 + (size_t)j * WT_SEL_CBINS1 : a.hist3 + (size_t)j * 2 * WT_SEL_CBINS);
}

__device__ __forceinline__ void clear_hist(Shared& sh, int n) {
    __syncthreads();
    for (int b = threadIdx.x; b < n; b += T) sh.hist[b] = 0u;
    __syncthreads();
}

// add the non-empty bins of the CTA's shared histogram to the job's global histogram
template <int PASS>
__device__ __forceinline__ void flush_hist(const WtSelArgs& a, Shared& sh, int j) {
    constexpr int NB = PASS == 1 ? WT_SEL_BINS1 : (PASS == 2 ? WT_SEL_BINS2 : WT_SEL_BINS3);
    constexpr int NBT = PASS == 1 ? NB : 2 * NB;
    unsigned long long* gh = global_hist<PASS>(a, j);
    __syncthreads();
    for (int b = threadIdx.x; b < NBT; b += T) {
        unsigned int v = sh.hist[b];
        if (v) atomicAdd(&gh[b], (unsigned long long)v);
        if (PASS == 1) {  // a warp covers 32 consecutive fine bins = one coarse bin
            unsigned int s = v;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
            if ((threadIdx.x & 31) == 0 && s) atomicAdd(&a.hist1c[(size_t)j * WT_SEL_COARSE + (b >> 5)], (unsigned long long)s);
        }
    }
}


// the same for nbt bins of the sampled path's candidate histograms
__device__ __forceinline__ void flush_bins(Shared& sh, unsigned long long* gh, int nbt) {
    __syncthreads();
    for (int b = threadIdx.x; b < nbt; b += T) {
        unsigned int v = sh.hist[b];
        if (v) atomicAdd(&gh[b], (unsigned long long)v);
    }
}

// Pass-1 lookup through the coarse histogram: warp 0 scans the 128 coarse sums (4 per lane), then warps 0
// and 1 scan the 32 fine bins of the coarse bin holding rank0 / rank1.  Two dependent L2 round trips and two
// barriers instead of a block-wide scan of 4096 bins.  Every thread returns the same result.
// n_pop > 0: the histogram is a SAMPLE of n_pop keys and rank0 <= rank1 are ranks in the population: they are
// scaled to the sample and pushed apart by 5 sigma of the binomial rank error (nothing when the sample is whole).
__device__ Found scan_find_two_level(const unsigned long long* __restrict__ coarse, const unsigned long long* __restrict__ fine,
                                     unsigned long long zeros, unsigned long long rank0, unsigned long long rank1, Shared& sh,
                                     long long n_pop = 0) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (wid == 0) {
        if (lane < 2) { sh.res_digit[lane] = 0u; sh.res_rem[lane] = 0ull; sh.res_cnt[lane] = 0ull; }  // empty histogram
        __syncwarp();
        unsigned long long c[4], s = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) c[i] = __ldcg(&coarse[lane * 4 + i]);
        if (lane == 0) c[0] += zeros;
#pragma unroll
        for (int i = 0; i < 4; ++i) s += c[i];
        unsigned long long inc = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        const unsigned long long total = __shfl_sync(0xffffffffu, inc, 31), excl = inc - s;
        unsigned long long rk[2] = {rank0, rank1};
        if (n_pop > 0 && total != (unsigned long long)n_pop) {
            // rank of a population quantile inside a sample of `total` keys: binomial, sigma^2 = total * q * (1 - q)
            const double f = (double)total / (double)n_pop;
            const double q0 = ((double)rank0 + 0.5) / (double)n_pop, q1 = ((double)rank1 + 0.5) / (double)n_pop;
            const double lo = floor((double)rank0 * f - 5.0 * sqrt((double)total * q0 * (1.0 - q0)) - 3.0);
            const double hi = ceil((double)rank1 * f + 5.0 * sqrt((double)total * q1 * (1.0 - q1)) + 3.0);
            rk[0] = lo > 0.0 ? (unsigned long long)lo : 0ull;
            rk[1] = hi > 0.0 ? (unsigned long long)hi : 0ull;
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            unsigned long long r = rk[q];
            if (r >= total) r = total ? total - 1 : 0;  // the padded sample rank may run past the sample
            if (s > 0 && excl <= r && r < excl + s) {
                unsigned long long run = excl;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (r < run + c[i]) { sh.res_digit[q] = lane * 4 + i; sh.res_rem[q] = r - run; break; }
                    run += c[i];
                }
            }
        }
    }
    __syncthreads();
    if (wid < 2) {
        const unsigned int cb = sh.res_digit[wid];
        const unsigned long long r = sh.res_rem[wid];
        const int b = (int)cb * 32 + lane;
        unsigned long long v = __ldcg(&fine[b]);
        if (b == 0) v += zeros;
        unsigned long long inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        const unsigned long long excl = inc - v;
        __syncwarp();  // every lane has read the coarse result before one of them overwrites it
        if (v > 0 && excl <= r && r < excl + v) {
            sh.res_digit[wid] = (unsigned int)b;
            sh.res_rem[wid] = r - excl;
            sh.res_cnt[wid] = v - (b == 0 ? zeros : 0ull);
        }
    }
    __syncthreads();
    Found f;
    f.d0 = sh.res_digit[0]; f.d1 = sh.res_digit[1];
    f.r0 = sh.res_rem[0]; f.r1 = sh.res_rem[1];
    f.c0 = sh.res_cnt[0]; f.c1 = sh.res_cnt[1];
    __syncthreads();  // the next call may overwrite the result slots
    return f;
}

// one call of f(j, job, base, len, tiles) per job overlapped by the CTA's tile range [t_begin, t_end)
template <class F>
__device__ __forceinline__ void for_each_segment(const WtSelArgs& a, int t_begin, int t_end, F f) {
    if (t_begin >= t_end) return;
    int j = (a.cta_job && a.cta_grid == (int)gridDim.x) ? __ldg(&a.cta_job[blockIdx.x]) : find_job(a.jobs, a.n_jobs, t_begin);
    int tile = t_begin;
    while (tile < t_end && j < a.n_jobs) {
        const WtSelJob job = a.jobs[j];
        if (tile >= job.tile_begin + job.n_tiles) { ++j; continue; }
        const int seg_end = min(t_end, job.tile_begin + job.n_tiles);
        const long long base = (long long)(tile - job.tile_begin) * WT_SEL_TILE;
        long long end = (long long)(seg_end - job.tile_begin) * WT_SEL_TILE;
        if (end > job.n) end = job.n;
        f(j, job, base, end - base, seg_end - tile);
        tile = seg_end;
        ++j;
    }
}


// calls op(key) on cand[0..n) (cand 16-byte aligned): uint4 loads from L2, UNROLL of them in flight per thread
template <class Op>
__device__ __forceinline__ void cand_for_each(const unsigned int* __restrict__ cand, unsigned int n, Op op) {
    const unsigned int nvec = n >> 2;
    const uint4* __restrict__ c4 = reinterpret_cast<const uint4*>(cand);
    for (unsigned int v = threadIdx.x; v < nvec; v += T * UNROLL) {
        uint4 r[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (v + u * T < nvec) r[u] = __ldcg(&c4[v + u * T]);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (v + u * T < nvec) { op(r[u].x); op(r[u].y); op(r[u].z); op(r[u].w); }
    }
    for (unsigned int i = (nvec << 2) + threadIdx.x; i < n; i += T) op(__ldcg(&cand[i]));
}


// ---- apply: out = |x| < t ? 0 : x, counts -----------------------------------------------------------------
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
        for (int w = 0; w < T / 32; ++w) { z += scratch[w]; nb += scratch[16 + w]; c += scratch[32 + w]; }
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


// P2 over one span: histogram of key bits 18..9 of the keys whose top digit is p0 / p1, and the same keys into
// the job's candidate list.  Staging is WARP-private (a slice of sh.stage per warp, fill level in a register,
// positions from a warp prefix sum), so the loop has no CTA barrier and warps overlap each other's load latency;
// a warp moves its slice to the global list (one global atomic) whenever it holds more than WSTAGE_FLUSH keys.
constexpr unsigned int WSTAGE_FLUSH = STAGE_FLUSH / (T / 32);
constexpr unsigned int WSTAGE = STAGE / (T / 32);  // = WSTAGE_FLUSH + 32 * UNROLL * 4: one more iteration always fits

__device__ __forceinline__ void warp_flush(const unsigned int* wbuf, unsigned int& wcount, unsigned int* g_cnt,
                                           unsigned int* __restrict__ cand, unsigned int cap) {
    const int lane = threadIdx.x & 31;
    unsigned int gbase = 0;
    if (lane == 0) gbase = atomicAdd(g_cnt, wcount);
    gbase = __shfl_sync(0xffffffffu, gbase, 0);
    __syncwarp();
    for (unsigned int i = lane; i < wcount; i += 32)
        if (gbase + i < cap) cand[gbase + i] = wbuf[i];
    __syncwarp();
    wcount = 0;
}

__device__ void hist2_compact_span(const float* x, long long len, unsigned int p0, unsigned int p1, bool two,
                                   Shared& sh, unsigned int* g_cnt, unsigned int* __restrict__ cand, unsigned int cap) {
    unsigned int* h = sh.hist;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned int* wbuf = sh.stage + wid * WSTAGE;
    unsigned int wcount = 0;  // warp-uniform
    auto matches = [&](unsigned int key) {
        const unsigned int top = key >> 19;
        return top == p0 || (two && top == p1);
    };
    auto keep = [&](unsigned int key, unsigned int pos) {
        atomicAdd(&h[((key >> 19) == p0 ? 0u : 1024u) + ((key >> 9) & 1023u)], 1u);
        wbuf[pos] = key;
    };
    long long done = 0;
    if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
        const long long nvec = len >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        for (long long v0 = 0; v0 < nvec; v0 += (long long)T * UNROLL) {  // CTA-uniform trip count
            const long long v = v0 + threadIdx.x;
            unsigned int key[UNROLL * 4];
            unsigned int hitmask = 0;
            float4 r[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
                if (v + (long long)u * T < nvec) r[u] = __ldcg(&x4[v + (long long)u * T]);
            if (WT_SEL_PREFETCH) {
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if (v + (long long)(UNROLL + u) * T < nvec) prefetch_l2(&x4[v + (long long)(UNROLL + u) * T]);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const bool ok = v + (long long)u * T < nvec;
                key[u * 4 + 0] = abs_key(r[u].x); key[u * 4 + 1] = abs_key(r[u].y);
                key[u * 4 + 2] = abs_key(r[u].z); key[u * 4 + 3] = abs_key(r[u].w);
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (ok && matches(key[u * 4 + c])) hitmask |= 1u << (u * 4 + c);
            }
            const unsigned int nh = __popc(hitmask);
            unsigned int inc = nh;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            const unsigned int total = __shfl_sync(0xffffffffu, inc, 31);
            if (total) {  // warp-uniform
                unsigned int pos = wcount + inc - nh;
#pragma unroll
                for (int e = 0; e < UNROLL * 4; ++e)
                    if (hitmask & (1u << e)) keep(key[e], pos++);
                wcount += total;
                if (wcount > WSTAGE_FLUSH) warp_flush(wbuf, wcount, g_cnt, cand, cap);
            }
        }
        done = nvec << 2;
    }
    for (long long i0 = done; i0 < len; i0 += T) {  // tail, or everything when x is not 16-byte aligned
        const bool ok = i0 + threadIdx.x < len;
        const unsigned int key = ok ? abs_key(__ldcg(&x[i0 + threadIdx.x])) : 0u;
        const bool hit = ok && matches(key);
        const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
        if (ballot) {
            if (hit) keep(key, wcount + __popc(ballot & ((1u << lane) - 1u)));
            wcount += __popc(ballot);
            if (wcount > WSTAGE_FLUSH) warp_flush(wbuf, wcount, g_cnt, cand, cap);
        }
    }
    if (wcount) warp_flush(wbuf, wcount, g_cnt, cand, cap);
}

// candidate-list boundary belonging to tile offset t of a job with n_tiles tiles (multiples of 4 keys)
__device__ __forceinline__ unsigned int cand_boundary(unsigned int n_c, int t, int n_tiles) {
    if (t >= n_tiles) return n_c;
    return (unsigned int)(((unsigned long long)n_c * (unsigned long long)t) / (unsigned long long)n_tiles) & ~3u;
}

// ================================ sampled path ===========================================================
// true when job j runs (or has been handed to) the radix path
__device__ __forceinline__ bool radix_job(const WtSelArgs& a, int j) { return a.path == 1 || __ldcg(&a.state[j].failed) != 0; }

// explicit-key ranks of the two order statistics: the implicit zeros sort first, so population rank k is
// explicit rank k - zeros (clamped: a rank inside the zeros needs no explicit key at all, see final_keys)
__device__ __forceinline__ void explicit_ranks(const WtSelJob& job, unsigned long long& e0, unsigned long long& e1) {
    const long long last = job.n - 1;
    long long r0 = job.k0 - job.zeros, r1 = job.k1 - job.zeros;
    r0 = r0 < 0 ? 0 : (r0 > last ? last : r0);
    r1 = r1 < 0 ? 0 : (r1 > last ? last : r1);
    if (r1 < r0) r1 = r0;
    e0 = (unsigned long long)r0;
    e1 = (unsigned long long)r1;
}

// A bracket of n_bins 12-bit bins spans 19 + ceil(log2(n_bins)) bits of key - lo_key: C1 resolves the top 12 of
// them, C2 the next min(cand_shift, 11), C3 -- only for brackets wider than 16 bins -- the last cand_shift2 <= 8.
__device__ __forceinline__ int cand_shift(unsigned int n_bins) { return 7 + (32 - __clz(n_bins - 1u)) * (n_bins > 1u); }
__device__ __forceinline__ int cand_shift2(int cs) { return cs > 11 ? cs - 11 : 0; }

// S0 over one segment: 12-bit histogram of the sample windows of `tiles` tiles starting at job tile `tile0`
__device__ __forceinline__ void sample_span(const float* x, long long len, int tile0, int tiles, int s,
                                            unsigned int* h) {
    if (s == 0) {
        span_for_each(x, len, [h](float v) { atomicAdd(&h[abs_key(v) >> 19], 1u); });
        return;
    }
    const int wshift = 12 - s;       // log2(window elements)
    const int vshift = wshift - 2;   // log2(float4s per window)
    const unsigned int wmask = (1u << s) - 1u, vmask = (1u << vshift) - 1u;
    const long long total = (long long)tiles << vshift;
    const bool aligned = (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
    for (long long v0 = threadIdx.x; v0 < total; v0 += (long long)T * UNROLL) {
        float4 r[UNROLL];
        bool ok[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const long long v = v0 + (long long)u * T;
            ok[u] = false;
            if (v < total) {
                const long long tile = v >> vshift;
                const unsigned int win = ((unsigned int)(tile0 + (int)tile) * 5u + 3u) & wmask;
                const long long off = (tile << 12) + ((long long)win << wshift) + (long long)(((unsigned int)v & vmask) << 2);
                if (off + 4 <= len) {
                    ok[u] = true;
                    if (aligned) r[u] = __ldcg(reinterpret_cast<const float4*>(x + off));
                    else r[u] = make_float4(__ldcg(x + off), __ldcg(x + off + 1), __ldcg(x + off + 2), __ldcg(x + off + 3));
                }
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            if (ok[u]) {
                atomicAdd(&h[abs_key(r[u].x) >> 19], 1u); atomicAdd(&h[abs_key(r[u].y) >> 19], 1u);
                atomicAdd(&h[abs_key(r[u].z) >> 19], 1u); atomicAdd(&h[abs_key(r[u].w) >> 19], 1u);
            }
        }
    }
}

// ---- S0: sample histogram -> hist1 / hist1c -------------------------------------------------------------
__device__ void phase_s0_sample(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        clear_hist(sh, WT_SEL_BINS1);
        acc.begin();
        sample_span(job.x + base, len, (int)(base / WT_SEL_TILE), tiles, job.sample_shift, sh.hist);
        acc.end(7);
        acc.begin();
        flush_hist<1>(a, sh, j);
        acc.end(8);
    });
}

// S1 over one span: count the keys below lo_key, compact the keys with key - lo_key < width into the job's
// candidate list.  Same warp-private staging as hist2_compact_span, and nothing else: no shared-memory atomics.
__device__ void range_compact_span(const float* x, long long len, unsigned int lo_key, unsigned int width,
                                   Shared& sh, unsigned long long* g_below, unsigned int* g_cnt,
                                   unsigned int* __restrict__ cand, unsigned int cap) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned int* wbuf = sh.stage + wid * WSTAGE;
    unsigned int wcount = 0;  // warp-uniform
    unsigned int nbelow = 0;
    long long done = 0;
    if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
        const long long nvec = len >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        for (long long v0 = 0; v0 < nvec; v0 += (long long)T * UNROLL) {  // CTA-uniform trip count
            const long long v = v0 + threadIdx.x;
            unsigned int key[UNROLL * 4];
            unsigned int hitmask = 0;
            float4 r[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
                if (v + (long long)u * T < nvec) r[u] = __ldcg(&x4[v + (long long)u * T]);
            if (WT_SEL_PREFETCH) {
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if (v + (long long)(UNROLL + u) * T < nvec) prefetch_l2(&x4[v + (long long)(UNROLL + u) * T]);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const bool ok = v + (long long)u * T < nvec;
                key[u * 4 + 0] = abs_key(r[u].x); key[u * 4 + 1] = abs_key(r[u].y);
                key[u * 4 + 2] = abs_key(r[u].z); key[u * 4 + 3] = abs_key(r[u].w);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if (ok) {
                        nbelow += key[u * 4 + c] < lo_key;
                        if (key[u * 4 + c] - lo_key < width) hitmask |= 1u << (u * 4 + c);
                    }
                }
            }
            if (__any_sync(0xffffffffu, hitmask != 0u)) {
                const unsigned int nh = __popc(hitmask);
                unsigned int inc = nh;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
                    if (lane >= d) inc += o;
                }
                const unsigned int total = __shfl_sync(0xffffffffu, inc, 31);
                unsigned int pos = wcount + inc - nh;
#pragma unroll
                for (int e = 0; e < UNROLL * 4; ++e)
                    if (hitmask & (1u << e)) wbuf[pos++] = key[e];
                wcount += total;
                if (wcount > WSTAGE_FLUSH) warp_flush(wbuf, wcount, g_cnt, cand, cap);
            }
        }
        done = nvec << 2;
    }
    for (long long i0 = done; i0 < len; i0 += T) {  // tail, or everything when x is not 16-byte aligned
        const bool ok = i0 + threadIdx.x < len;
        const unsigned int key = ok ? abs_key(__ldcg(&x[i0 + threadIdx.x])) : 0u;
        nbelow += ok && key < lo_key;
        const bool hit = ok && (key - lo_key < width);
        const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
        if (ballot) {
            if (hit) wbuf[wcount + __popc(ballot & ((1u << lane) - 1u))] = key;
            wcount += __popc(ballot);
            if (wcount > WSTAGE_FLUSH) warp_flush(wbuf, wcount, g_cnt, cand, cap);
        }
    }
    if (wcount) warp_flush(wbuf, wcount, g_cnt, cand, cap);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) nbelow += __shfl_down_sync(0xffffffffu, nbelow, d);
    if (lane == 0 && nbelow) atomicAdd(g_below, (unsigned long long)nbelow);
}

// ---- S1: bracket from the sample histogram; the one full pass: count below, compact the bracket -------------
__device__ void phase_s1_compact(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        WtSelState* st = &a.state[j];
        unsigned long long e0, e1;
        explicit_ranks(job, e0, e1);
        acc.begin();
        const Found f = scan_find_two_level(a.hist1c + (size_t)j * WT_SEL_COARSE, global_hist<1>(a, j), 0ull, e0, e1, sh, job.n);
        acc.end(9);
        const unsigned int lo_bin = f.d0;
        unsigned int n_bins = f.d1 - f.d0 + 1u;
        if (f.d1 < f.d0 || job.cand == nullptr) n_bins = 0u;  // -> radix path
        if (base == 0 && threadIdx.x == 0) {
            st->lo_bin = lo_bin; st->n_bins = n_bins;
            if (n_bins > 16u) *a.wide_flag = 1;
        }
        if (n_bins == 0u) return;
        acc.begin();
        // a bracket that starts at bin 0 leaves the exact zeros out of the list: they are counted as "below" and a
        // rank inside them is 0 without looking (the massive tie of this domain: pruned weights, vanishing details)
        const unsigned int lo_eff = (lo_bin << 19) | (lo_bin == 0u ? 1u : 0u);
        range_compact_span(job.x + base, len, lo_eff, ((lo_bin + n_bins) << 19) - lo_eff, sh, &st->below, &st->s_count, job.cand,
                           job.cand_cap);
        acc.end(10);
    });
}

// ---- T: one CTA finishes one job: verify the bracket, then an MSD radix select (11-bit digits, both ranks at once)
// over the job's candidate list, which is short (5 sigma of the sampling error: ~1e4 keys) and L2-resident.  CTA-local
// barriers only -- the grid-wide candidate passes C1..C3 cost a grid barrier, a flush and 296 redundant scans each
// and are kept for the lists this CTA would take too long over (big_flag).
__device__ void phase_tail(const WtSelArgs& a, Shared& sh, unsigned int tail_max) {
    for (int j = blockIdx.x; j < a.n_jobs; j += gridDim.x) {
        const WtSelJob job = a.jobs[j];
        WtSelState* st = &a.state[j];
        const unsigned int n_bins = __ldcg(&st->n_bins), n_c = __ldcg(&st->s_count);
        const unsigned long long below = __ldcg(&st->below);
        unsigned long long e0, e1;
        explicit_ranks(job, e0, e1);
        const bool zlo = __ldcg(&st->lo_bin) == 0u;  // then `below` counts the exact zeros
        const bool in0 = (zlo && e0 < below) || (below <= e0 && e0 < below + (unsigned long long)n_c);
        const bool in1 = (zlo && e1 < below) || (below <= e1 && e1 < below + (unsigned long long)n_c);
        const bool good = n_bins != 0u && n_c <= job.cand_cap && in0 && in1;
        if (!good) {  // hist1 / hist1c held the sample: the radix path starts from zero
            unsigned long long* gh1 = global_hist<1>(a, j);
            for (int b = threadIdx.x; b < WT_SEL_BINS1; b += T) gh1[b] = 0ull;
            if (threadIdx.x < WT_SEL_COARSE) a.hist1c[(size_t)j * WT_SEL_COARSE + threadIdx.x] = 0ull;
            if (threadIdx.x == 0) { st->failed = 1; *a.fail_flag = 1; }
            continue;
        }
        if (n_c > tail_max) {
            if (threadIdx.x == 0) *a.big_flag = 1;
            continue;
        }
        const unsigned int lo_key = __ldcg(&st->lo_bin) << 19;
        unsigned int pre0 = 0u, pre1 = 0u;                                   // digits resolved so far of key - lo_key
        unsigned long long r0 = e0 < below ? 0ull : e0 - below, r1 = e1 < below ? 0ull : e1 - below;  // ranks inside them
        int prev = 19 + (n_bins > 1u ? 32 - __clz(n_bins - 1u) : 0);          // bits of key - lo_key still open
        unsigned int* h = sh.hist;                                           // two histograms of 2048 bins
        while (prev > 0) {
            const int s = prev > 11 ? prev - 11 : 0;
            const unsigned int mask = (1u << (prev - s)) - 1u;
            const bool two = pre0 != pre1;
            clear_hist(sh, 4096);
            cand_for_each(job.cand, n_c, [=](unsigned int k) {
                const unsigned int v = k - lo_key, top = v >> prev;
                if (top == pre0) atomicAdd(&h[(v >> s) & mask], 1u);
                else if (two && top == pre1) atomicAdd(&h[2048u + ((v >> s) & mask)], 1u);
            });
            __syncthreads();
            const Found f = scan_find_pair<2048, false>(h, two ? h + 2048 : h, 0ull, 0ull, r0, r1, sh);
            pre0 = (pre0 << (prev - s)) | f.d0;
            pre1 = (pre1 << (prev - s)) | f.d1;
            r0 = f.r0;
            r1 = f.r1;
            prev = s;
        }
        // both order statistics and the threshold, the way final_keys / P4 form them
        const unsigned long long ez = lo_key == 0u ? below : 0ull;  // exact zeros among the explicit keys
        const unsigned int key0 = (job.k0 < job.zeros || e0 < ez) ? 0u : lo_key + pre0;
        const unsigned int key1 = (job.k1 < job.zeros || e1 < ez) ? 0u : lo_key + pre1;
        const float t = lerp_threshold(__uint_as_float(key0), __uint_as_float(key1), job.gamma, job.lerp_mode);
        if (threadIdx.x == 0) {
            st->a = __uint_as_float(key0);
            st->b = __uint_as_float(key1);
            st->t = t;
            a.thresholds[job.layer] = t;
        }
        if (threadIdx.x == 0) {
            st->tail_key[0] = pre0;
            st->tail_key[1] = pre1;
            st->tail_done = 1;
        }
    }
}

// ---- C1: verify the bracket; histogram of the top 12 bits of key - lo_key over the candidate list -> hist2 ------
__device__ void phase_c1_cand(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc, unsigned int tail_max) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        WtSelState* st = &a.state[j];
        const unsigned int n_bins = __ldcg(&st->n_bins), n_c = __ldcg(&st->s_count);
        if (n_c <= tail_max) return;  // the one-CTA tail's job (it runs in this same phase: its flags are not read here)
        const unsigned long long below = __ldcg(&st->below);
        unsigned long long e0, e1;
        explicit_ranks(job, e0, e1);
        const bool zlo = __ldcg(&st->lo_bin) == 0u;  // then `below` counts the exact zeros
        const bool in0 = (zlo && e0 < below) || (below <= e0 && e0 < below + (unsigned long long)n_c);
        const bool in1 = (zlo && e1 < below) || (below <= e1 && e1 < below + (unsigned long long)n_c);
        const bool good = n_bins != 0u && n_c <= job.cand_cap && in0 && in1;
        if (!good) {  // every CTA of the job sees the same counts and takes the same decision
            if (base == 0) {  // hist1 / hist1c held the sample: the radix path starts from zero
                unsigned long long* gh1 = global_hist<1>(a, j);
                for (int b = threadIdx.x; b < WT_SEL_BINS1; b += T) gh1[b] = 0ull;
                if (threadIdx.x < WT_SEL_COARSE) a.hist1c[(size_t)j * WT_SEL_COARSE + threadIdx.x] = 0ull;
                if (threadIdx.x == 0) { st->failed = 1; *a.fail_flag = 1; }
            }
            return;
        }
        acc.begin();
        if (n_bins > 16u && base == 0) {  // C3 reuses the slot of the sample histogram, dead since the last barrier
            unsigned long long* gh1 = global_hist<1>(a, j);
            for (int b = threadIdx.x; b < 2 * WT_SEL_CBINS3; b += T) gh1[b] = 0ull;
        }
        clear_hist(sh, WT_SEL_CBINS1);
        unsigned int* h = sh.hist;
        const unsigned int lo_key = __ldcg(&st->lo_bin) << 19;
        const int cs = cand_shift(n_bins);
        const int t0 = (int)(base / WT_SEL_TILE);
        const unsigned int lo = cand_boundary(n_c, t0, job.n_tiles), hi = cand_boundary(n_c, t0 + tiles, job.n_tiles);
        cand_for_each(job.cand + lo, hi - lo, [h, lo_key, cs](unsigned int k) { atomicAdd(&h[(k - lo_key) >> cs], 1u); });
        flush_bins(sh, global_hist<2>(a, j), WT_SEL_CBINS1);
        acc.end(11);
    });
}

// ---- C2: digit of both ranks from hist2; histogram of the low bits of the matching candidates -> hist3 ---------
__device__ void phase_c2_cand(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        WtSelState* st = &a.state[j];
        if (__ldcg(&st->failed) || __ldcg(&st->tail_done)) return;
        const unsigned int n_c = __ldcg(&st->s_count);
        const unsigned long long below = __ldcg(&st->below);
        unsigned long long e0, e1;
        explicit_ranks(job, e0, e1);
        acc.begin();
        const unsigned long long* gh = global_hist<2>(a, j);
        // a rank inside the exact zeros (bracket from bin 0 only) needs no candidate: final_keys answers 0
        const Found f = scan_find_pair<WT_SEL_CBINS1, true>(gh, gh, 0ull, 0ull, e0 < below ? 0ull : e0 - below,
                                                            e1 < below ? 0ull : e1 - below, sh);
        const unsigned int d0 = f.d0, d1 = f.d1;
        const bool two = (d0 != d1);
        if (base == 0 && threadIdx.x == 0) {
            st->cprefix[0] = d0; st->cprefix[1] = d1;
            st->crem[0] = f.r0; st->crem[1] = f.r1;
        }
        clear_hist(sh, 2 * WT_SEL_CBINS);
        unsigned int* h = sh.hist;
        const unsigned int lo_key = __ldcg(&st->lo_bin) << 19;
        const int t0 = (int)(base / WT_SEL_TILE);
        const unsigned int lo = cand_boundary(n_c, t0, job.n_tiles), hi = cand_boundary(n_c, t0 + tiles, job.n_tiles);
        const int cs = cand_shift(__ldcg(&st->n_bins)), cs2 = cand_shift2(cs);
        const unsigned int low = (1u << (cs - cs2)) - 1u;
        cand_for_each(job.cand + lo, hi - lo, [h, lo_key, d0, d1, two, cs, cs2, low](unsigned int k) {
            const unsigned int kk = k - lo_key, top = kk >> cs;
            if (top == d0) atomicAdd(&h[(kk >> cs2) & low], 1u);
            else if (two && top == d1) atomicAdd(&h[WT_SEL_CBINS + ((kk >> cs2) & low)], 1u);
        });
        flush_bins(sh, a.hist3 + (size_t)j * 2 * WT_SEL_CBINS, 2 * WT_SEL_CBINS);
        acc.end(12);
    });
}

// ---- C3 (only when some bracket is wider than 16 bins): digits from hist3; histogram of the last <= 8 bits ------
__device__ void phase_c3_cand(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        WtSelState* st = &a.state[j];
        const unsigned int n_bins = __ldcg(&st->n_bins);
        if (__ldcg(&st->failed) || __ldcg(&st->tail_done) || n_bins <= 16u) return;
        const unsigned int n_c = __ldcg(&st->s_count);
        const unsigned int d0 = __ldcg(&st->cprefix[0]), d1 = __ldcg(&st->cprefix[1]);
        const unsigned long long r0 = __ldcg(&st->crem[0]), r1 = __ldcg(&st->crem[1]);
        const unsigned long long* gh3 = global_hist<3>(a, j);
        const Found f = scan_find_pair<WT_SEL_CBINS, false>(gh3, d0 != d1 ? gh3 + WT_SEL_CBINS : gh3, 0ull, 0ull, r0, r1, sh);
        const unsigned int q0 = (d0 << 11) | f.d0, q1 = (d1 << 11) | f.d1;  // wide: C2 resolved exactly 11 bits
        const bool two = (q0 != q1);
        if (base == 0 && threadIdx.x == 0) {
            st->cprefix2[0] = q0; st->cprefix2[1] = q1;
            st->crem2[0] = f.r0; st->crem2[1] = f.r1;
        }
        clear_hist(sh, 2 * WT_SEL_CBINS3);
        unsigned int* h = sh.hist;
        const unsigned int lo_key = __ldcg(&st->lo_bin) << 19;
        const int cs2 = cand_shift2(cand_shift(n_bins));
        const unsigned int low = (1u << cs2) - 1u;
        const int t0 = (int)(base / WT_SEL_TILE);
        const unsigned int lo = cand_boundary(n_c, t0, job.n_tiles), hi = cand_boundary(n_c, t0 + tiles, job.n_tiles);
        cand_for_each(job.cand + lo, hi - lo, [h, lo_key, q0, q1, two, cs2, low](unsigned int k) {
            const unsigned int kk = k - lo_key, top = kk >> cs2;
            if (top == q0) atomicAdd(&h[kk & low], 1u);
            else if (two && top == q1) atomicAdd(&h[WT_SEL_CBINS3 + (kk & low)], 1u);
        });
        flush_bins(sh, global_hist<1>(a, j), 2 * WT_SEL_CBINS3);
    });
}

// both order statistics of job j as key patterns, from whichever path finished the job
__device__ __forceinline__ void final_keys(const WtSelArgs& a, Shared& sh, int j, const WtSelJob& job, unsigned int& key0,
                                           unsigned int& key1) {
    WtSelState* st = &a.state[j];
    if (radix_job(a, j)) {
        const unsigned int n0 = __ldcg(&st->prefix2[0]), n1 = __ldcg(&st->prefix2[1]);
        const unsigned long long r0 = __ldcg(&st->rem2[0]), r1 = __ldcg(&st->rem2[1]);
        const bool two2 = (n0 != n1);
        const unsigned long long zeros = (unsigned long long)job.zeros;
        const unsigned long long* gh3 = global_hist<3>(a, j);
        const Found f = scan_find_pair<WT_SEL_BINS3, false>(gh3, two2 ? gh3 + WT_SEL_BINS3 : gh3, n0 == 0u ? zeros : 0ull,
                                                     n1 == 0u ? zeros : 0ull, r0, r1, sh);
        key0 = (n0 << 9) | f.d0;
        key1 = (n1 << 9) | f.d1;
    } else {
        const unsigned int lo_key = __ldcg(&st->lo_bin) << 19;
        const unsigned int n_bins = __ldcg(&st->n_bins);
        const int cs = cand_shift(n_bins), cs2 = cand_shift2(cs);
        unsigned int v0, v1;  // key - lo_key
        if (__ldcg(&st->tail_done)) {
            v0 = __ldcg(&st->tail_key[0]);
            v1 = __ldcg(&st->tail_key[1]);
        } else if (n_bins > 16u) {
            const unsigned int q0 = __ldcg(&st->cprefix2[0]), q1 = __ldcg(&st->cprefix2[1]);
            const unsigned long long r0 = __ldcg(&st->crem2[0]), r1 = __ldcg(&st->crem2[1]);
            const unsigned long long* gh = global_hist<1>(a, j);
            const Found f = scan_find_pair<WT_SEL_CBINS3, false>(gh, q0 != q1 ? gh + WT_SEL_CBINS3 : gh, 0ull, 0ull, r0, r1, sh);
            v0 = (q0 << cs2) | f.d0;
            v1 = (q1 << cs2) | f.d1;
        } else {
            const unsigned int d0 = __ldcg(&st->cprefix[0]), d1 = __ldcg(&st->cprefix[1]);
            const unsigned long long r0 = __ldcg(&st->crem[0]), r1 = __ldcg(&st->crem[1]);
            const unsigned long long* gh3 = global_hist<3>(a, j);
            const Found f = scan_find_pair<WT_SEL_CBINS, false>(gh3, d0 != d1 ? gh3 + WT_SEL_CBINS : gh3, 0ull, 0ull, r0, r1, sh);
            v0 = (d0 << cs) | f.d0;
            v1 = (d1 << cs) | f.d1;
        }
        unsigned long long e0, e1;
        explicit_ranks(job, e0, e1);
        const unsigned long long ez = lo_key == 0u ? __ldcg(&st->below) : 0ull;  // exact zeros among the explicit keys
        key0 = (job.k0 < job.zeros || e0 < ez) ? 0u : lo_key + v0;  // a rank inside the zeros is 0
        key1 = (job.k1 < job.zeros || e1 < ez) ? 0u : lo_key + v1;
    }
}

// ================================ radix path ================================================================
// ---- P1: histogram of key bits 30..19 ------------------------------------------------------------------
__device__ void phase1_hist(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        if (!radix_job(a, j)) return;
        clear_hist(sh, WT_SEL_BINS1);
        unsigned int* h = sh.hist;
        span_for_each(job.x + base, len, [h](float v) { atomicAdd(&h[abs_key(v) >> 19], 1u); });
        flush_hist<1>(a, sh, j);
    });
}

// ---- P2: digit of both ranks from hist1; histogram of bits 18..9 + compaction of the matching keys ------
__device__ void phase2_hist_compact(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        if (!radix_job(a, j)) return;
        WtSelState* st = &a.state[j];
        const unsigned long long* gh1 = global_hist<1>(a, j);
        const Found f = scan_find_two_level(a.hist1c + (size_t)j * WT_SEL_COARSE, gh1, (unsigned long long)job.zeros,
                                            (unsigned long long)job.k0, (unsigned long long)job.k1, sh);
        const unsigned int p0 = f.d0, p1 = f.d1;
        const bool two = (p0 != p1);
        // compaction only when every key of the one or two selected digits fits the candidate list
        const unsigned long long ncand = f.c0 + (two ? f.c1 : 0ull);
        const bool ovf = (job.cand == nullptr) || (ncand > (unsigned long long)job.cand_cap);
        if (base == 0 && threadIdx.x == 0) {  // the CTA holding the job's first tile publishes for P3
            st->prefix1[0] = p0; st->prefix1[1] = p1;
            st->rem1[0] = f.r0; st->rem1[1] = f.r1;
            st->overflow = ovf ? 1 : 0;
        }
        clear_hist(sh, 2 * WT_SEL_BINS2);
        unsigned int* h = sh.hist;
        if (ovf) {
            span_for_each(job.x + base, len, [h, p0, p1, two](float v) { bin_one<2>(abs_key(v), h, p0, p1, two); });
        } else {
            hist2_compact_span(job.x + base, len, p0, p1, two, sh, &st->cand_count, job.cand, job.cand_cap);
        }
        flush_hist<2>(a, sh, j);
    });
}

// ---- P3: 22-bit prefixes from hist2; histogram of bits 8..0 over the candidate list ----------------------
__device__ void phase3_hist_cand(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        if (!radix_job(a, j)) return;
        WtSelState* st = &a.state[j];
        const unsigned int p0 = __ldcg(&st->prefix1[0]), p1 = __ldcg(&st->prefix1[1]);
        const unsigned long long r0 = __ldcg(&st->rem1[0]), r1 = __ldcg(&st->rem1[1]);
        const bool ovf = __ldcg(&st->overflow) != 0;
        const bool two = (p0 != p1);
        const unsigned long long zeros = (unsigned long long)job.zeros;
        const unsigned long long* gh2 = global_hist<2>(a, j);
        const Found f = scan_find_pair<WT_SEL_BINS2, false>(gh2, two ? gh2 + WT_SEL_BINS2 : gh2, p0 == 0u ? zeros : 0ull,
                                                     p1 == 0u ? zeros : 0ull, r0, r1, sh);
        const unsigned int n0 = (p0 << 10) | f.d0, n1 = (p1 << 10) | f.d1;
        const bool two2 = (n0 != n1);
        if (base == 0 && threadIdx.x == 0) {
            st->prefix2[0] = n0; st->prefix2[1] = n1;
            st->rem2[0] = f.r0; st->rem2[1] = f.r1;
        }
        clear_hist(sh, 2 * WT_SEL_BINS3);
        unsigned int* h = sh.hist;
        if (ovf) {
            span_for_each(job.x + base, len, [h, n0, n1, two2](float v) { bin_one<3>(abs_key(v), h, n0, n1, two2); });
        } else {
            unsigned int n_c = __ldcg(&st->cand_count);
            if (n_c > job.cand_cap) n_c = job.cand_cap;  // cannot happen: the overflow test uses exact counts
            const int t0 = (int)(base / WT_SEL_TILE);
            const unsigned int lo = cand_boundary(n_c, t0, job.n_tiles), hi = cand_boundary(n_c, t0 + tiles, job.n_tiles);
            cand_for_each(job.cand + lo, hi - lo, [h, n0, n1, two2](unsigned int k) { bin_one<3>(k, h, n0, n1, two2); });
        }
        flush_hist<3>(a, sh, j);
    });
}

// stores o[0..len) at the same offset of every peer's buffer; the thread -> element mapping is the apply loop's
__device__ __noinline__ void mirror_span(const float* o, long long len, const WtPeers* __restrict__ peers) {
    const int np = __ldg(&peers->n);
    long long done = 0;
    if (__ldg(&peers->mc)) {   // NVLS: one store per value to the multicast mapping
        char* g = reinterpret_cast<char*>(const_cast<float*>(o)) + __ldg(&peers->delta[0]);
        if ((reinterpret_cast<uintptr_t>(o) & 15u) == 0) {
            const long long nvec = len >> 2;
            for (long long v = threadIdx.x; v < nvec; v += T) wt_mc_store4(g + v * 16, reinterpret_cast<const float4*>(o)[v]);
            done = nvec << 2;
        }
        for (long long i = done + threadIdx.x; i < len; i += T) wt_mc_store1(g + i * 4, o[i]);
        return;
    }
    if ((reinterpret_cast<uintptr_t>(o) & 15u) == 0) {
        const long long nvec = len >> 2;
        const float4* o4 = reinterpret_cast<const float4*>(o);
        for (long long v = threadIdx.x; v < nvec; v += T) {
            const float4 r = o4[v];
            for (int q = 0; q < np; ++q)
                __stcs(reinterpret_cast<float4*>(reinterpret_cast<char*>(const_cast<float4*>(&o4[v])) + __ldg(&peers->delta[q])), r);
        }
        done = nvec << 2;
    }
    for (long long i = done + threadIdx.x; i < len; i += T) {
        const float w = o[i];
        for (int q = 0; q < np; ++q) *reinterpret_cast<float*>(reinterpret_cast<char*>(const_cast<float*>(&o[i])) + __ldg(&peers->delta[q])) = w;
    }
}

// ---- P4: threshold from hist3; apply: out = |x| < t ? 0 : x, counts ----------------------------------------
__device__ void phase4_apply(const WtSelArgs& a, Shared& sh, int t_begin, int t_end, Acc& acc) {
    for_each_segment(a, t_begin, t_end, [&](int j, const WtSelJob& job, long long base, long long len, int tiles) {
        if (!job.out && base != 0) return;  // select-only job: its first-tile CTA publishes, nobody applies
        WtSelState* st = &a.state[j];
        if (!job.out && __ldcg(&st->tail_done)) return;  // the one-CTA tail already published a, b and the threshold
        acc.begin();
        unsigned int key0, key1;
        final_keys(a, sh, j, job, key0, key1);
        const float t = lerp_threshold(__uint_as_float(key0), __uint_as_float(key1), job.gamma, job.lerp_mode);
        if (base == 0 && threadIdx.x == 0) {
            st->a = __uint_as_float(key0);
            st->b = __uint_as_float(key1);
            st->t = t;
            a.thresholds[job.layer] = t;
            if (job.out) a.counts[(size_t)job.layer * 4] = job.n;
        }
        acc.end(13);
        if (!job.out) return;
        acc.begin();
        const float* x = job.x + base;  // may alias out (in-place pruning): no __restrict__, no read-only path
        float* o = job.out + base;
        unsigned int zc = 0, nzb = 0;
        long long done = 0;
        if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(o)) & 15u) == 0) {
            const long long nvec = len >> 2;
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float4* o4 = reinterpret_cast<float4*>(o);
            for (long long v = threadIdx.x; v < nvec; v += (long long)T * UNROLL) {
                float4 r[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if (v + (long long)u * T < nvec) r[u] = __ldcg(&x4[v + (long long)u * T]);
                if (WT_SEL_PREFETCH) {
#pragma unroll
                    for (int u = 0; u < UNROLL; ++u)
                        if (v + (long long)(UNROLL + u) * T < nvec) prefetch_l2(&x4[v + (long long)(UNROLL + u) * T]);
                }
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    if (v + (long long)u * T < nvec) {
                        r[u].x = apply_one(r[u].x, t, zc, nzb); r[u].y = apply_one(r[u].y, t, zc, nzb);
                        r[u].z = apply_one(r[u].z, t, zc, nzb); r[u].w = apply_one(r[u].w, t, zc, nzb);
                        __stcs(&o4[v + (long long)u * T], r[u]);
                    }
                }
            }
            done = nvec << 2;
        }
        for (long long i = done + threadIdx.x; i < len; i += T) o[i] = apply_one(x[i], t, zc, nzb);
        // layer-sharded run: the span just written (every thread re-reads exactly what it stored, still in L2) also goes
        // to the same slot of every peer's buffer; out of line so the single-GPU loop keeps its registers
        // (the apply loop may have taken its scalar path -- x misaligned -- while mirror_span vectorises on o's alignment
        // alone: then a thread re-reads elements other threads stored, so the CTA meets here first)
        if (__ldg(&a.jobs[j].peers.n)) {
            __syncthreads();
            mirror_span(o, len, &a.jobs[j].peers);
        }
        // every element of the span is visited exactly once: thread 0 contributes the span length
        block_add3(a.counts + (size_t)job.layer * 4, zc, nzb, threadIdx.x == 0 ? (unsigned int)len : 0u, sh.hist);
        acc.end(14);
    });
}

__global__ void __launch_bounds__(WT_SEL_THREADS, 2) wt_select_kernel(WtSelArgs a, int do_apply) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Shared& sh = *reinterpret_cast<Shared*>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const long long NT = a.n_tiles;
    int t_begin, t_end;
    if (a.cta_tiles && a.cta_grid == (int)gridDim.x) {  // cost-balanced split built by the plan
        t_begin = __ldg(&a.cta_tiles[blockIdx.x]);
        t_end = __ldg(&a.cta_tiles[blockIdx.x + 1]);
    } else {
        t_begin = (int)((NT * blockIdx.x) / gridDim.x);
        t_end = (int)((NT * (blockIdx.x + 1)) / gridDim.x);
    }
    Acc acc;
    acc.slot = a.stamps ? a.stamps + (size_t)blockIdx.x * 16 : nullptr;
    acc.t0 = 0;
    acc.stamp(0);
    if (a.zero_next) {   // the scratch + counts block of the plan's next run (in place of a memset in front of every step)
        int4* z = reinterpret_cast<int4*>(a.zero_next);
        const unsigned long long n16 = a.zero_next_bytes >> 4;
        for (unsigned long long i = (unsigned long long)blockIdx.x * T + threadIdx.x; i < n16; i += (unsigned long long)gridDim.x * T)
            z[i] = make_int4(0, 0, 0, 0);
    }
    if (a.path != 1) {
        if (a.path == 0) {   // path 2 (precompacted): the analysis kernels did S0 and S1 while they held the coefficients
            phase_s0_sample(a, sh, t_begin, t_end, acc);
            acc.stamp(1);
            grid.sync();
            phase_s1_compact(a, sh, t_begin, t_end, acc);
            acc.stamp(2);
            grid.sync();
        }
        // one phase: the short lists are finished by one CTA each (tail), the long ones (s_count > tail_max, final since the
        // last barrier) get their first-digit histogram from everybody (C1) meanwhile -- round 1 ran the tail, a barrier, then
        // C1: at p = 90 the headline's three 146 k-key lists made every step pay both, one after the other
        // How long a list the tail takes depends on whether the grid passes run anyway: when no list exceeds tail_max, every
        // job ends in its tail and C2 / C3 never start (p = 50: 179 us against 200); when one does, C2 will run, and lists
        // over WT_SEL_TAIL_SHORT go with it rather than keep one CTA busy for three passes (p = 90: 195 us against 201).
        // The counts are final since the last barrier, so every CTA takes the same decision.
        unsigned int tail_max = a.tail_max;
        {
            bool big = false;
            for (int j = threadIdx.x; j < a.n_jobs; j += T) big |= __ldcg(&a.state[j].s_count) > a.tail_max;
            if (__syncthreads_or(big) && tail_max > WT_SEL_TAIL_SHORT) tail_max = WT_SEL_TAIL_SHORT;
        }
        phase_tail(a, sh, tail_max);
        acc.stamp(3);
        phase_c1_cand(a, sh, t_begin, t_end, acc, tail_max);
        grid.sync();
        if (__ldcg(a.big_flag) != 0) {  // grid-uniform (final before the last barrier): some list is too long for one CTA
            phase_c2_cand(a, sh, t_begin, t_end, acc);
            acc.stamp(4);
            grid.sync();
            if (__ldcg(a.wide_flag) != 0) {  // grid-uniform: set during S1
                phase_c3_cand(a, sh, t_begin, t_end, acc);
                grid.sync();
            }
        }
    }
    if (a.path == 1 || __ldcg(a.fail_flag) != 0) {  // grid-uniform: the flag was final before the last barrier
        acc.begin();
        phase1_hist(a, sh, t_begin, t_end, acc);
        if (a.path == 1) acc.stamp(1);  // radix-only launches reuse the sampled path's boundary slots: arrival at the
        grid.sync();                    // barrier after P1 / P2 / P3 (mean vs max over CTAs = imbalance)
        phase2_hist_compact(a, sh, t_begin, t_end, acc);
        if (a.path == 1) acc.stamp(2);
        grid.sync();
        phase3_hist_cand(a, sh, t_begin, t_end, acc);
        if (a.path == 1) acc.stamp(3);
        grid.sync();
        acc.end(15);
    }
    acc.stamp(5);
    phase4_apply(a, sh, t_begin, t_end, acc);
    acc.stamp(6);
}

// Precompacted launches: one CTA per job turns the job's SAMPLE histogram (written by the analysis kernels' sample pass)
// into the bracket the analysis kernels then compact -- exactly what every CTA of phase_s1_compact works out for itself.
// Jobs the analysis kernels do not cover (sample_shift < 0: level-0 tensors, generic-path layers) are handed to the radix
// passes of the same select launch.
__global__ void __launch_bounds__(WT_SEL_THREADS) wt_select_bracket_kernel(WtSelArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Shared& sh = *reinterpret_cast<Shared*>(smem_raw);
    const int j = blockIdx.x;
    const WtSelJob job = a.jobs[j];
    WtSelState* st = &a.state[j];
    if (job.sample_shift < 0 || job.cand == nullptr) {
        if (threadIdx.x == 0) { st->failed = 1; st->n_bins = 0u; *a.fail_flag = 1; }
        return;
    }
    unsigned long long e0, e1;
    explicit_ranks(job, e0, e1);
    const Found f = scan_find_two_level(a.hist1c + (size_t)j * WT_SEL_COARSE, global_hist<1>(a, j), 0ull, e0, e1, sh, job.n);
    unsigned int n_bins = f.d1 - f.d0 + 1u;
    if (f.d1 < f.d0) n_bins = 0u;
    if (threadIdx.x == 0) {
        st->lo_bin = f.d0;
        st->n_bins = n_bins;
        if (n_bins > 16u) *a.wide_flag = 1;
    }
}

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
    size_t per = (size_t)WT_SEL_BINS1 * 8 + (size_t)WT_SEL_CBINS1 * 8 + 2 * (size_t)WT_SEL_CBINS * 8 +
                 (size_t)WT_SEL_COARSE * 8 + sizeof(WtSelState);
    return per * (size_t)n_jobs + 256;
}

void wt_select_rebind(WtSelArgs* a, void* scratch, int n_jobs) {
    char* p = static_cast<char*>(scratch);
    a->hist1 = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * WT_SEL_BINS1 * 8;
    a->hist2 = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * WT_SEL_CBINS1 * 8;
    a->hist3 = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * 2 * WT_SEL_CBINS * 8;
    a->hist1c = reinterpret_cast<unsigned long long*>(p); p += (size_t)n_jobs * WT_SEL_COARSE * 8;
    a->state = reinterpret_cast<WtSelState*>(p); p += (size_t)n_jobs * sizeof(WtSelState);
    a->fail_flag = reinterpret_cast<int*>(p);  // inside the 256 spare bytes
    a->wide_flag = a->fail_flag + 1;
    a->big_flag = a->fail_flag + 2;
}

void wt_select_bind(WtSelArgs* a, void* scratch, int n_jobs) {
    wt_select_rebind(a, scratch, n_jobs);
    a->zero_next = nullptr;
    a->zero_next_bytes = 0;
    a->path = -1;  // resolved at launch: wt_select_path()
    a->short_lists = 0;
    a->stamps = nullptr;  // tuning aid, set by the plan when WT_SELECT_STAMPS is in the environment
    a->cta_tiles = nullptr;
    a->cta_job = nullptr;
    a->cta_grid = 0;
    a->n_jobs = n_jobs;
}

static int select_blocks_per_sm(const WtDeviceInfo& di, int* nb_out) {
    static int blocks_per_sm[64] = {0};  // per device; racing threads compute the same value
    WT_CHECK(di.coop, WT_E_UNSUPPORTED, "device does not support cooperative launch");
    WT_CHECK(di.device >= 0 && di.device < 64, WT_E_UNSUPPORTED, "device ordinal %d out of range", di.device);
    if (blocks_per_sm[di.device] == 0) {
        WT_CUDA(cudaFuncSetAttribute(wt_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Shared)));
        int nb = 0;
        WT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wt_select_kernel, WT_SEL_THREADS, sizeof(Shared)));
        WT_CHECK(nb > 0, WT_E_CUDA, "select kernel does not fit on an SM");
        blocks_per_sm[di.device] = nb;
    }
    *nb_out = blocks_per_sm[di.device];
    return WT_OK;
}

int wt_select_grid(int n_tiles, int* grid_out) {
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    int nb = 0;
    if ((rc = select_blocks_per_sm(di, &nb))) return rc;
    int grid = nb * di.sm_count;
    if (grid > n_tiles) grid = n_tiles;
    if (grid < 1) grid = 1;
    *grid_out = grid;
    return WT_OK;
}

// Contiguous split of the tile space that minimises the largest CTA cost, cost = tiles + seg_cost per segment (a
// segment = the piece of one job inside one CTA: every phase pays a scan, a histogram clear and a flush for it).
// Greedy packing under a capacity, smallest feasible capacity by bisection.
static int pack_tiles(const WtSelJob* jobs, int n_jobs, int grid, long long cap, int seg_cost, int* tiles_out, int* job_out) {
    int cta = 0;
    long long room = cap;
    if (tiles_out) { tiles_out[0] = 0; job_out[0] = 0; }
    for (int j = 0; j < n_jobs; ++j) {
        int left = jobs[j].n_tiles, tile = jobs[j].tile_begin;
        while (left > 0) {
            if (room <= seg_cost) {  // not even one tile of a new segment fits: next CTA
                ++cta;
                room = cap;
                if (tiles_out && cta < grid) { tiles_out[cta] = tile; job_out[cta] = j; }
                if (cta >= grid && !tiles_out) return cta + 1;
            }
            long long take = room - seg_cost;
            if (take > left) take = left;
            room -= seg_cost + take;
            left -= (int)take;
            tile += (int)take;
        }
    }
    return cta + 1;  // CTAs used
}

void wt_select_balance(const WtSelJob* jobs, int n_jobs, int n_tiles, int grid, int* tiles_out, int* job_out) {
    static const int seg_cost = [] { const char* e = getenv("WT_SEL_SEGMENT_COST"); return e ? atoi(e) : WT_SEL_SEGMENT_COST; }();
    long long lo = seg_cost + 1, hi = (long long)n_tiles + (long long)seg_cost * n_jobs + 1;
    while (lo < hi) {
        const long long mid = (lo + hi) / 2;
        if (pack_tiles(jobs, n_jobs, grid, mid, seg_cost, nullptr, nullptr) <= grid) hi = mid; else lo = mid + 1;
    }
    const int used = pack_tiles(jobs, n_jobs, grid, lo, seg_cost, tiles_out, job_out);
    for (int c = used; c <= grid; ++c) tiles_out[c] = n_tiles;  // unused CTAs get empty ranges
    for (int c = used; c < grid; ++c) job_out[c] = n_jobs > 0 ? n_jobs - 1 : 0;
}

// Which path a launch over n_tiles tiles takes.  Measured on one B200 (tools/select_probe.py, tools/bank_probe.py,
// tools/fused_probe.py): the sampled path saves one of the two full passes but runs one or two more phases, each worth
// 6..8 us of barrier, scan and flush latency whatever the size.  What decides is whether the keys stay in the 126 MB L2
// between the radix path's two passes: 79 MB of haar coefficients or 45 MB of raw weights do (radix wins: 138 vs 144 us,
// 63 vs 81 us), 100 MB (ResNet-18 bior2.2 L3) and 124 MB (bior4.4 L5) do not (sampled wins: 152 vs 165 us, 205 vs 212 us,
// long candidate lists included).  WT_SELECT_PATH=radix|sampled (read at every launch) overrides.
int wt_select_path(int n_tiles, int short_lists) {
    const char* e = getenv("WT_SELECT_PATH");
    if (e && !strcmp(e, "radix")) return 1;
    if (e && !strcmp(e, "sampled")) return 0;
    return (n_tiles >= WT_SEL_SAMPLED_MIN_TILES || short_lists) ? 0 : 1;
}

// Below 40 Mi keys the sampled path only pays when the one-CTA tail can finish every job (measured, ResNet-18: 81 us
// against 104 us radix when it can, 83 us against 64 us when the lists are long and the grid-wide passes C1..C3 run).
// A candidate list is the bracket of 12-bit bins around the rank: one bin (2^-4 of an octave) holds about q / 16 of
// the explicit keys at explicit quantile q of a bell-shaped magnitude distribution (density of |x| per log-bin grows
// linearly with x near zero), and the bracket adds 5 sigma of the sample's rank error on both sides.  Deep
// decompositions of small planes put the rank a few percent into the explicit keys (padding zeros and vanishing
// deep-level details sort first): short lists.  A mid-quantile over raw weights: long lists, radix.  A wrong guess only
// costs time: the kernel checks the real list length (big_flag) and the real bracket (fail_flag).
int wt_select_predict_short(const WtSelJob* jobs, int n_jobs) {
    static const double limit = [] { const char* e = getenv("WT_SEL_SHORT_LIMIT"); return e ? atof(e) : 32768.0; }();
    for (int j = 0; j < n_jobs; ++j) {
        const WtSelJob& jb = jobs[j];
        if (jb.n <= 0) continue;
        if (jb.cand == nullptr) return 0;
        double r = (double)(jb.k0 - jb.zeros);
        if (r < 0) r = 0;
        double q = r / (double)jb.n;
        if (q > 1.0) q = 1.0;
        const double m = (double)(jb.n >> jb.sample_shift);
        const double est = (double)jb.n * (2.0 * q / 16.0 + (m < (double)jb.n ? 10.0 * sqrt(q * (1.0 - q) / m) : 0.0));
        if (est > limit) return 0;
    }
    return 1;
}

int wt_select_launch(const WtSelArgs& args, cudaStream_t stream) {
    int grid = 1;
    int rc = wt_select_grid(args.n_tiles, &grid);
    if (rc) return rc;
    const size_t smem = sizeof(Shared);
    WtSelArgs a = args;
    a.path = a.precompacted ? 2 : wt_select_path(a.n_tiles, a.short_lists);
    {
        const char* e = getenv("WT_SEL_TAIL_MAX");  // tuning aid, read at every launch
        a.tail_max = e ? (unsigned int)atoi(e) : (unsigned int)WT_SEL_TAIL_MAX;
    }
    int do_apply = a.any_apply;  // jobs without `out` skip the apply loop inside P4
    void* params[] = {&a, &do_apply};
    WT_CUDA(cudaLaunchCooperativeKernel((const void*)wt_select_kernel, dim3(grid), dim3(WT_SEL_THREADS), params, smem, stream));
    return WT_OK;
}

int wt_select_bracket_launch(const WtSelArgs& args, cudaStream_t stream) {
    if (args.n_jobs <= 0) return WT_OK;
    static bool attr_set[64] = {false};
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    if (di.device >= 0 && di.device < 64 && !attr_set[di.device]) {
        WT_CUDA(cudaFuncSetAttribute(wt_select_bracket_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Shared)));
        attr_set[di.device] = true;
    }
    wt_select_bracket_kernel<<<args.n_jobs, WT_SEL_THREADS, sizeof(Shared), stream>>>(args);
    WT_CUDA(cudaGetLastError());
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

