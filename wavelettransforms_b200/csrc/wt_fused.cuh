// wt_fused.cuh -- the fused prune kernel: ONE persistent launch per plane class does the whole
//   wavedec2 -> np.percentile(|c|, p) -> np.where(|c| < t, 0, c) -> waverec2 -> crop -> zero count
// of /root/reference/ResNet/dwt_pruning.py:35-95 (and the absolute-threshold flow of
// dwt_pruning_NoEntropy.py:12-62) for every tensor of the class, WITHOUT ever writing a coefficient to HBM: the
// coefficients of a small plane live in registers and are recomputed from the weights in every pass that needs them
// (hard part 2).  Algorithmic traffic stays what it is -- read the weight, write the pruned weight.
//
// One cooperative launch per plane class, a persistent grid (every CTA owns one contiguous range of the class's items,
// cut into per-job segments) and a few grid barriers:
//   S   12-bit histogram (key bits 30..19 of |c|) of a 1/2^s SAMPLE of every job's keys (whole for jobs < 32 Ki keys)
//   --- barrier ---
//   C   every CTA scans the sample histogram of its segment's job -> a bracket of consecutive bins that holds both
//       ranks unless the sample lied (5 sigma of the binomial rank error); then the one full pass: count the keys
//       below the bracket, compact the keys inside it into the job's candidate list (warp-private staging) and
//       histogram the top 11 bits of (key - lo) of those candidates (F1)
//   --- barrier ---
//   T1  the candidate lists, cut into slices dealt to all CTAs: verify the bracket with the EXACT counts (a miss hands
//       the job to the exact path X below, so the result never depends on the sample), digits of both ranks from F1,
//       histogram of the next <= 11 bits of the matching candidates (F2)
//   --- barrier ---
//   T2  only when some bracket is wider than 8 bins (a rank next to the massive tie at zero): the last <= 9 bits (F3)
//   A   every CTA scans F2 (F3) of its segment's job -> sorted[k0], sorted[k1] -> np.percentile's lerp (legacy_1_22 /
//       nep50_2x) -> threshold; apply: recompute the coefficients, zero |c| < t, synthesise, crop, store, count zeros
// No phase has a serial tail: after a barrier every CTA scans the small L2-resident histogram of its own job
// redundantly instead of waiting for an elected CTA; the CTA that holds a job's first item (slice) publishes the
// result of the scan for the phase after the next barrier, so every phase does one scan per segment.
//
// X1..X3 + XA: exact MSD radix select (12 + 11 + 8 bits, three recomputing passes, no candidate list) for the jobs
// whose bracket missed or whose list overflowed; four more phases over the same partition, run only if some job failed.
//
// Absolute-threshold classes (dwt_pruning_NoEntropy.py:45) run phase A alone: one pass, no barrier.
//
// The per-run scratch (per-job state and histograms) must be zero at launch; every launch zeroes the scratch block of
// the NEXT run (two blocks alternate), so a step enqueues no memset.  (A first version of this kernel replaced the grid
// barriers by per-job flags and a ticket queue; measured on the B200 every hand-over between stages of a job cost
// 10-25 us under load and the waiting CTAs could do nothing useful: 150 us for ResNet-18 at level 0 against 66 us for
// the barrier kernel it was meant to beat.  Dropped.)
#pragma once
#include <cooperative_groups.h>
#include <math.h>

#include "wt_common.cuh"

namespace wtf {

constexpr int NB1 = WT_F_BINS1;     // key bits 30..19
constexpr int NCOARSE = NB1 / 32;
constexpr int NBF = WT_F_FBINS;     // F1 / F2 / X2 digits: 11 bits
constexpr int NBF3 = WT_F_FBINS3;   // F3: <= 9 bits, X3: 8 bits

__device__ __forceinline__ unsigned int abs_key(float v) { return __float_as_uint(v) & 0x7fffffffu; }

__device__ __forceinline__ unsigned int ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP) + mbarrier ---------------------------------------------------
__device__ __forceinline__ unsigned int smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
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
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, unsigned int bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- np.percentile's lerp ------------------------------------------------------------------------------
__device__ __forceinline__ float lerp_threshold(float a, float b, double gamma, int mode) {
    if (mode == WT_LERP_LEGACY) {
        // numpy 1.22 _lerp on 0-d operands: (b - a) in float32, the rest in float64, then the float32 array comparison
        // demotes the float64 scalar.  No FMA.
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

// explicit-key ranks of the two order statistics: the implicit zeros sort first, so population rank k is explicit
// rank k - zeros (clamped: a rank inside the zeros needs no explicit key at all, see final_threshold)
__device__ __forceinline__ void explicit_ranks(const WtFJob& job, unsigned long long& e0, unsigned long long& e1) {
    const long long last = job.n - 1;
    long long r0 = job.k0 - job.zeros, r1 = job.k1 - job.zeros;
    r0 = r0 < 0 ? 0 : (r0 > last ? last : r0);
    r1 = r1 < 0 ? 0 : (r1 > last ? last : r1);
    if (r1 < r0) r1 = r0;
    e0 = (unsigned long long)r0;
    e1 = (unsigned long long)r1;
}

struct Found {
    unsigned int d0, d1;        // digit holding rank0 / rank1
    unsigned long long r0, r1;  // rank inside that digit
    unsigned long long c0, c1;  // EXPLICIT keys in that digit (implicit zeros excluded)
};

template <int T>
struct Shared {
    unsigned int hist[NB1];               // S / X1: 4096 bins; C: F1 [0, 2048); F2 / X2: 2 x 2048; F3 / X3: 2 x 512
    unsigned int stage[T / 32 * WT_F_WSTAGE];  // C: candidate keys on their way to the global list, one slice per warp
    unsigned long long warp_sums[2][T / 32 + 1];
    unsigned long long res_rem[2];
    unsigned long long res_cnt[2];
    unsigned int res_digit[2];
    unsigned int red[3][T / 32];
    int jidx[2][WT_F_JIDX];   // first item of every job in the full / sample item space
    WtFJob job;       // the job of the current segment
};

// ---- block scans over global (L2) histograms -------------------------------------------------------------------------
// Locates rank0 in h0 (+ z0 implicit zeros in bin 0) and rank1 in h1 (+ z1); h1 == h0 when both ranks share a histogram.
// Two passes over the bins (sum, then the owning thread walks its bins again) keep the register footprint flat.
// Every thread returns the same result.
template <int T, int NB>
__device__ Found scan_find_pair(const unsigned int* h0, const unsigned int* h1, unsigned long long z0, unsigned long long z1,
                                unsigned long long rank0, unsigned long long rank1, Shared<T>& sh) {
    constexpr int PER = (NB + T - 1) / T;
    const int b0 = threadIdx.x * PER;
    unsigned long long s0 = 0, s1 = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int b = b0 + i;
        if (b < NB) {
            s0 += __ldcg(&h0[b]);
            s1 += __ldcg(&h1[b]);
        }
    }
    if (threadIdx.x == 0) { s0 += z0; s1 += z1; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long i0 = s0, i1 = s1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned long long o0 = __shfl_up_sync(0xffffffffu, i0, d), o1 = __shfl_up_sync(0xffffffffu, i1, d);
        if (lane >= d) { i0 += o0; i1 += o1; }
    }
    if (threadIdx.x < 2) { sh.res_digit[threadIdx.x] = 0u; sh.res_rem[threadIdx.x] = 0ull; sh.res_cnt[threadIdx.x] = 0ull; }
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
        for (int i = 0; i < PER; ++i) {
            const int b = b0 + i;
            if (b >= NB) break;
            const unsigned long long c = (unsigned long long)__ldcg(&h0[b]) + (b == 0 ? z0 : 0ull);
            if (rank0 < run + c) {
                sh.res_digit[0] = b;
                sh.res_rem[0] = rank0 - run;
                sh.res_cnt[0] = c - (b == 0 ? z0 : 0ull);
                break;
            }
            run += c;
        }
    }
    if (s1 > 0 && e1 <= rank1 && rank1 < e1 + s1) {
        unsigned long long run = e1;
        for (int i = 0; i < PER; ++i) {
            const int b = b0 + i;
            if (b >= NB) break;
            const unsigned long long c = (unsigned long long)__ldcg(&h1[b]) + (b == 0 ? z1 : 0ull);
            if (rank1 < run + c) {
                sh.res_digit[1] = b;
                sh.res_rem[1] = rank1 - run;
                sh.res_cnt[1] = c - (b == 0 ? z1 : 0ull);
                break;
            }
            run += c;
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

// 12-bit lookup through the coarse histogram: warp 0 scans the 128 coarse sums (4 per lane), then warps 0 and 1 scan
// the 32 fine bins of the coarse bin holding rank0 / rank1: two dependent L2 round trips instead of a scan of 4096
// bins.  n_pop > 0: the histogram is a SAMPLE of a population of n_pop keys and rank0 <= rank1 are ranks in the
// population: they are scaled to the sample and pushed apart by 5 sigma of the binomial rank error (nothing when the
// sample is whole).  Every thread returns the same result.
template <int T>
__device__ Found scan_find_two_level(const unsigned int* coarse, const unsigned int* fine, unsigned long long zeros,
                                     unsigned long long rank0, unsigned long long rank1, Shared<T>& sh, long long n_pop) {
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
    __syncthreads();
    return f;
}

// candidate-list boundary of slice i of n (multiples of 4 keys)
__device__ __forceinline__ unsigned int cand_boundary(unsigned int n_c, int i, int n) {
    if (i >= n) return n_c;
    return (unsigned int)(((unsigned long long)n_c * (unsigned long long)i) / (unsigned long long)n) & ~3u;
}

// calls op(key) on cand[0..n) (cand 16-byte aligned): uint4 loads from L2, 4 of them in flight per thread
template <int T, class Op>
__device__ __forceinline__ void cand_for_each(const unsigned int* cand, unsigned int n, Op op) {
    const unsigned int nvec = n >> 2;
    const uint4* c4 = reinterpret_cast<const uint4*>(cand);
    for (unsigned int v = threadIdx.x; v < nvec; v += T * 4) {
        uint4 r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (v + u * T < nvec) r[u] = __ldcg(&c4[v + u * T]);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (v + u * T < nvec) { op(r[u].x); op(r[u].y); op(r[u].z); op(r[u].w); }
    }
    for (unsigned int i = (nvec << 2) + threadIdx.x; i < n; i += T) op(__ldcg(&cand[i]));
}

// tuning aid (WT_FUSED_STAMPS): one globaltimer stamp per phase boundary and CTA
__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// the histogram passes over N keys per thread: S / X1 bin key bits 30..19, X2 bits 18..8 and X3 bits 7..0 of the keys
// that match the prefix found so far
enum { HIST_TOP = 0, HIST_X2 = 1, HIST_X3 = 2 };
struct HistOp {
    unsigned int* h;
    int kind;
    unsigned int p0, p1;
    bool two;
    template <int N>
    __device__ __forceinline__ void operator()(const unsigned int (&key)[N]) const {
        if (kind == HIST_TOP) {
#pragma unroll
            for (int i = 0; i < N; ++i)
                if (key[i] != WT_F_NOKEY) atomicAdd(&h[key[i] >> 19], 1u);
        } else if (kind == HIST_X2) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                const unsigned int top = key[i] >> 19;   // WT_F_NOKEY >> 19 = 8191 is no 12-bit prefix
                if (top == p0) atomicAdd(&h[(key[i] >> 8) & 2047u], 1u);
                else if (two && top == p1) atomicAdd(&h[NBF + ((key[i] >> 8) & 2047u)], 1u);
            }
        } else {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                const unsigned int top = key[i] >> 8;
                if (top == p0) atomicAdd(&h[key[i] & 255u], 1u);
                else if (two && top == p1) atomicAdd(&h[NBF3 + (key[i] & 255u)], 1u);
            }
        }
    }
};

// moves a warp's staged candidate keys to the job's global list (one global atomic per flush)
__device__ __forceinline__ void warp_flush(const unsigned int* wbuf, unsigned int& wc, unsigned int* g_count, unsigned int* cand,
                                           unsigned int cap) {
    const int lane = threadIdx.x & 31;
    unsigned int gbase = 0;
    if (lane == 0) gbase = atomicAdd(g_count, wc);
    gbase = __shfl_sync(0xffffffffu, gbase, 0);
    __syncwarp();
    for (unsigned int i = lane; i < wc; i += 32)
        if (gbase + i < cap) cand[gbase + i] = wbuf[i];
    __syncwarp();
    wc = 0;
}

// the count / compact pass (phase C) over N keys per thread: keys below the bracket are counted, keys inside it go to
// the warp's staging slice (positions from one warp prefix sum per group of <= WT_F_GROUP keys) and into the F1 histogram
struct CompactOp {
    unsigned int* h;
    unsigned int* wbuf;
    unsigned int* g_count;
    unsigned int* cand;
    unsigned int cap;
    unsigned int lo, w;
    int s1;
    unsigned int nb, wc;   // hot accumulators, read back by the caller
    template <int N>
    __device__ __forceinline__ void operator()(const unsigned int (&key)[N]) {
#pragma unroll
        for (int g0 = 0; g0 < N; g0 += WT_F_GROUP) {
            constexpr int G = WT_F_GROUP;
            unsigned int hitmask = 0;
#pragma unroll
            for (int i = 0; i < G; ++i) {
                if (g0 + i < N) {
                    nb += key[g0 + i] < lo;
                    if (key[g0 + i] - lo < w) hitmask |= 1u << i;
                }
            }
            if (__any_sync(0xffffffffu, hitmask != 0u)) {
                const int lane = threadIdx.x & 31;
                const unsigned int nh = __popc(hitmask);
                unsigned int inc = nh;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
                    if (lane >= d) inc += o;
                }
                const unsigned int total = __shfl_sync(0xffffffffu, inc, 31);
                unsigned int pos = wc + inc - nh;
#pragma unroll
                for (int i = 0; i < G; ++i) {
                    if (g0 + i < N && (hitmask & (1u << i))) {
                        wbuf[pos++] = key[g0 + i];
                        atomicAdd(&h[(key[g0 + i] - lo) >> s1], 1u);
                    }
                }
                wc += total;
                if (wc > WT_F_WFLUSH) warp_flush(wbuf, wc, g_count, cand, cap);
            }
        }
    }
};

// ======================================================================================================================
// The CTA-side phases.  K (the plane class) provides:
//   THREADS, MIN_CTAS, SMEM_BYTES (class-private shared memory after Shared<THREADS>), init(smem), numel(job)
//   visit(bk, job, first, count, sampled, smem, op)   op(keys[N]) over every explicit key of items [first, first + count)
//                                                     (sampled: of the sample items), N keys per thread at a time
//                                                     (WT_F_NOKEY = none), called by all 32 lanes of a warp together
//   apply(bk, job, first, count, t, smem, zc, nzb)    out = prune(in) for the items; zc += zeros written, nzb += nonzero inputs
// ======================================================================================================================
template <class K>
struct Cta {
    static constexpr int T = K::THREADS;
    using Sh = Shared<T>;

    const WtBank& bk;
    const WtFArgs& a;
    Sh& sh;
    unsigned char* ksm;  // class-private shared memory

    __device__ Cta(const WtBank& bk_, const WtFArgs& a_, Sh& sh_, unsigned char* ksm_) : bk(bk_), a(a_), sh(sh_), ksm(ksm_) {}

    __device__ __forceinline__ WtFHeader* header() const { return reinterpret_cast<WtFHeader*>(a.scratch); }
    __device__ __forceinline__ WtFState* state(int j) const {
        return reinterpret_cast<WtFState*>(a.scratch + sizeof(WtFHeader)) + j;
    }
    __device__ __forceinline__ unsigned int* hists(int j) const {
        return reinterpret_cast<unsigned int*>(a.scratch + sizeof(WtFHeader) + (size_t)a.n_jobs * sizeof(WtFState)) +
               (size_t)j * WT_F_HIST_WORDS;
    }
    __device__ __forceinline__ unsigned int* hist1(int j) const { return hists(j); }
    __device__ __forceinline__ unsigned int* coarse(int j) const { return hists(j) + NB1; }
    __device__ __forceinline__ unsigned int* histf1(int j) const { return hists(j) + NB1 + NCOARSE; }
    __device__ __forceinline__ unsigned int* histf2(int j) const { return hists(j) + NB1 + NCOARSE + NBF; }
    __device__ __forceinline__ unsigned int* histf3(int j) const { return hists(j) + NB1 + NCOARSE + 3 * NBF; }

    __device__ __forceinline__ void clear_hist(int n) {
        __syncthreads();
        for (int b = threadIdx.x; b < n; b += T) sh.hist[b] = 0u;
        __syncthreads();
    }
    __device__ __forceinline__ void flush_bins(unsigned int* gh, int n) {
        __syncthreads();
        for (int b = threadIdx.x; b < n; b += T) {
            const unsigned int v = sh.hist[b];
            if (v) atomicAdd(&gh[b], v);
        }
    }
    __device__ __forceinline__ void flush_hist1(int j) {
        unsigned int* gh = hist1(j);
        unsigned int* gc = coarse(j);
        __syncthreads();
        for (int b = threadIdx.x; b < NB1; b += T) {  // a warp covers 32 consecutive fine bins = one coarse bin
            const unsigned int v = sh.hist[b];
            if (v) atomicAdd(&gh[b], v);
            unsigned int s = v;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
            if ((threadIdx.x & 31) == 0 && s) atomicAdd(&gc[b >> 5], s);
        }
    }
    // the job descriptor into shared memory (int4 copies)
    __device__ __forceinline__ void load_job(int j) {
        __syncthreads();
        const int4* src = reinterpret_cast<const int4*>(a.jobs + j);
        int4* dst = reinterpret_cast<int4*>(&sh.job);
        for (int i = threadIdx.x; i < (int)(sizeof(WtFJob) / 16); i += T) dst[i] = __ldg(&src[i]);
        __syncthreads();
    }
    __device__ __forceinline__ int find_job(int item, bool sample) const {
        int lo = 0, hi = a.n_jobs - 1;
        const bool cached = a.n_jobs <= WT_F_JIDX;   // offsets in shared memory: no chain of dependent global loads
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            const int b = cached ? sh.jidx[sample ? 1 : 0][mid] : (sample ? __ldg(&a.jobs[mid].s_begin) : __ldg(&a.jobs[mid].item_begin));
            if (b <= item) lo = mid; else hi = mid - 1;
        }
        return lo;
    }
    // f(j, first, count) for every job overlapped by this CTA's contiguous share of the class's (sample) items; the
    // job's descriptor is in sh.job when f runs
    template <class F>
    __device__ __forceinline__ void for_each_segment(bool sample, F f) {
        const long long total = sample ? a.total_s_items : a.total_items;
        const int lo = (int)(total * blockIdx.x / gridDim.x), hi = (int)(total * (blockIdx.x + 1) / gridDim.x);
        if (lo >= hi) return;
        int j = find_job(lo, sample), item = lo;
        while (item < hi && j < a.n_jobs) {
            load_job(j);
            const int jb = sample ? sh.job.s_begin : sh.job.item_begin, jn = sample ? sh.job.s_items : sh.job.items;
            if (item < jb + jn) {
                const int end = hi < jb + jn ? hi : jb + jn;
                f(j, item - jb, end - item);
                item = end;
            }
            ++j;
        }
    }

    // bits of (key - lo) below the F1 digit, taken by the F2 digit, left for F3
    __device__ __forceinline__ static void tail_shifts(unsigned int w, int& s1, int& take2, int& s2) {
        const int bits = 32 - __clz(w - 1u);
        s1 = bits - 11;
        take2 = s1 < 11 ? s1 : 11;
        s2 = s1 - take2;
    }

    // ---- S / X1: 12-bit histogram of the sample (S) or of every key of the failed jobs (X1) ------------------------------
    __device__ __noinline__ void phase_hist_top(bool sample) {
        for_each_segment(sample, [&](int j, int first, int count) {
            if (!sample && !__ldcg(&state(j)->failed)) return;
            if (sample && sh.job.lerp_mode == WT_LERP_ABSOLUTE) return;
            clear_hist(NB1);
            HistOp op{sh.hist, HIST_TOP, 0u, 0u, false};
            K::visit(bk, sh.job, first, count, sample, ksm, op);
            flush_hist1(j);
        });
    }

    __device__ void mark_failed(WtFState* st) {   // one thread
        st->failed = 1;
        atomicAdd(&header()->fail_count, 1u);
    }

    // ---- C: bracket from the sample histogram; count below, compact the bracket, F1 histogram -----------------------------
    __device__ __noinline__ void phase_compact() {
        for_each_segment(false, [&](int j, int first, int count) {
            const WtFJob& job = sh.job;
            if (job.lerp_mode == WT_LERP_ABSOLUTE) return;
            WtFState* st = state(j);
            unsigned long long e0, e1;
            explicit_ranks(job, e0, e1);
            const Found f = scan_find_two_level<T>(coarse(j), hist1(j), 0ull, e0, e1, sh, job.n);
            const bool ok = !(f.d1 < f.d0 || job.cand == nullptr);
            // a bracket that starts at bin 0 leaves the exact zeros out of the list: they are counted as "below" and a
            // rank inside them is 0 without looking (the massive tie of this domain: pruned weights, vanishing details)
            const unsigned int lo = (f.d0 << 19) | (f.d0 == 0u ? 1u : 0u);
            const unsigned int w = ((f.d1 + 1u) << 19) - lo;
            if (first == 0 && threadIdx.x == 0) {   // every CTA of the job takes the same decision from the same histogram
                st->lo_key = lo;
                st->width = ok ? w : 0u;
                if (!ok) mark_failed(st);
                else if (w > (1u << 22)) header()->wide = 1u;
            }
            if (!ok) return;
            int s1, take2, s2;
            tail_shifts(w, s1, take2, s2);
            clear_hist(NBF);
            CompactOp op{sh.hist, sh.stage + (threadIdx.x >> 5) * WT_F_WSTAGE, &st->s_count, job.cand, job.cand_cap, lo, w, s1, 0u, 0u};
            K::visit(bk, job, first, count, false, ksm, op);
            if (op.wc) warp_flush(op.wbuf, op.wc, op.g_count, op.cand, op.cap);
            unsigned int nb = op.nb;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) nb += __shfl_down_sync(0xffffffffu, nb, d);
            if ((threadIdx.x & 31) == 0 && nb) atomicAdd(&st->below, (unsigned long long)nb);
            flush_bins(histf1(j), NBF);
        });
    }

    // exact-count check of the bracket C used: both ranks fall inside the candidate list (or inside the exact zeros below it)
    __device__ __forceinline__ bool bracket_holds(const WtFJob& job, unsigned int lo, unsigned int n_c, unsigned long long below,
                                                  unsigned long long& e0, unsigned long long& e1) const {
        explicit_ranks(job, e0, e1);
        const bool zlo = lo == 1u;  // the bracket starts at bin 0: `below` counts the exact zeros
        const bool in0 = (zlo && e0 < below) || (below <= e0 && e0 < below + (unsigned long long)n_c);
        const bool in1 = (zlo && e1 < below) || (below <= e1 && e1 < below + (unsigned long long)n_c);
        return n_c <= job.cand_cap && in0 && in1;
    }

    // ---- T1 / T2: the candidate lists in slices of WT_F_SLICE keys; slice sl of job j belongs to CTA (37 j + sl) mod grid ----
    __device__ __noinline__ void phase_tail(int round) {
        unsigned int* cnt = sh.stage;   // free outside phase C: every job's list length, read once
        __syncthreads();
        for (int i = threadIdx.x; i < a.n_jobs; i += T) cnt[i] = __ldcg(&state(i)->s_count);
        __syncthreads();
        const unsigned int G = gridDim.x;
        for (int j = 0; j < a.n_jobs; ++j) {
            unsigned int n_c = cnt[j];
            const unsigned int slices = n_c ? (n_c + WT_F_SLICE - 1) / WT_F_SLICE : 1u;
            const unsigned int sl0 = (blockIdx.x + G - (unsigned int)((37ull * j) % G)) % G;
            if (sl0 >= slices) continue;   // CTA-uniform
            load_job(j);
            const WtFJob& job = sh.job;
            WtFState* st = state(j);
            unsigned int* gh1 = hist1(j);
            const bool owner = sl0 == 0u;
            const unsigned int lo = __ldcg(&st->lo_key), w = __ldcg(&st->width);
            if (__ldcg(&st->failed) || w == 0u) {   // failed in C (or, round 2, in T1): the sample histogram must not reach X1
                if (owner && round == 1) for (int b = threadIdx.x; b < NB1 + NCOARSE; b += T) gh1[b] = 0u;
                continue;
            }
            const unsigned long long below = __ldcg(&st->below);
            unsigned long long e0, e1;
            if (!bracket_holds(job, lo, n_c, below, e0, e1)) {   // every CTA of the job sees the same counts
                if (owner && round == 1) {
                    for (int b = threadIdx.x; b < NB1 + NCOARSE; b += T) gh1[b] = 0u;
                    if (threadIdx.x == 0) mark_failed(st);
                }
                continue;
            }
            int s1, take2, s2;
            tail_shifts(w, s1, take2, s2);
            unsigned int p0, p1;
            unsigned int* h = sh.hist;
            if (round == 1) {
                const unsigned int* g1 = histf1(j);
                const Found f = scan_find_pair<T, NBF>(g1, g1, 0ull, 0ull, e0 < below ? 0ull : e0 - below, e1 < below ? 0ull : e1 - below, sh);
                p0 = f.d0; p1 = f.d1;
                if (owner && threadIdx.x == 0) { st->d[0] = p0; st->d[1] = p1; st->r[0] = f.r0; st->r[1] = f.r1; }
                if (n_c == 0u) continue;
                clear_hist(2 * NBF);
            } else {
                if (s2 == 0) continue;
                const unsigned int d0 = __ldcg(&st->d[0]), d1 = __ldcg(&st->d[1]);
                const unsigned int* g2 = histf2(j);
                const Found f = scan_find_pair<T, NBF>(g2, d0 != d1 ? g2 + NBF : g2, 0ull, 0ull, __ldcg(&st->r[0]), __ldcg(&st->r[1]), sh);
                p0 = (d0 << take2) | f.d0; p1 = (d1 << take2) | f.d1;
                if (owner && threadIdx.x == 0) { st->d2[0] = p0; st->d2[1] = p1; st->r2[0] = f.r0; st->r2[1] = f.r1; }
                if (n_c == 0u) continue;
                clear_hist(2 * NBF3);
            }
            const bool tw = p0 != p1;
            for (unsigned int sl = sl0; sl < slices; sl += G) {
                const unsigned int c0 = sl * WT_F_SLICE, len = n_c - c0 < WT_F_SLICE ? n_c - c0 : WT_F_SLICE;
                if (round == 1) {
                    const unsigned int mask = (1u << take2) - 1u;
                    cand_for_each<T>(job.cand + c0, len, [=](unsigned int k) {
                        const unsigned int v = k - lo, top = v >> s1;
                        if (top == p0) atomicAdd(&h[(v >> s2) & mask], 1u);
                        else if (tw && top == p1) atomicAdd(&h[NBF + ((v >> s2) & mask)], 1u);
                    });
                } else {
                    const unsigned int mask = (1u << s2) - 1u;
                    cand_for_each<T>(job.cand + c0, len, [=](unsigned int k) {
                        const unsigned int v = k - lo, top = v >> s2;
                        if (top == p0) atomicAdd(&h[v & mask], 1u);
                        else if (tw && top == p1) atomicAdd(&h[NBF3 + (v & mask)], 1u);
                    });
                }
            }
            if (round == 1) flush_bins(histf2(j), 2 * NBF); else flush_bins(histf3(j), 2 * NBF3);
        }
    }

    // ---- X2 / X3: the exact path's second and third digit over every key of the failed jobs -------------------------------
    __device__ __noinline__ void phase_exact(int which) {
        for_each_segment(false, [&](int j, int first, int count) {
            WtFState* st = state(j);
            if (!__ldcg(&st->failed)) return;
            const WtFJob& job = sh.job;
            const unsigned long long z = (unsigned long long)job.zeros;
            unsigned int p0, p1;
            if (which == 2) {
                const Found f = scan_find_two_level<T>(coarse(j), hist1(j), z, (unsigned long long)job.k0, (unsigned long long)job.k1, sh, 0);
                p0 = f.d0; p1 = f.d1;
                if (first == 0 && threadIdx.x == 0) { st->d[0] = p0; st->d[1] = p1; st->r[0] = f.r0; st->r[1] = f.r1; }
                clear_hist(2 * NBF);
            } else {
                const unsigned int d0 = __ldcg(&st->d[0]), d1 = __ldcg(&st->d[1]);
                const unsigned int* g2 = histf2(j);
                const Found f = scan_find_pair<T, NBF>(g2, d0 != d1 ? g2 + NBF : g2, d0 == 0u ? z : 0ull, d1 == 0u ? z : 0ull,
                                                       __ldcg(&st->r[0]), __ldcg(&st->r[1]), sh);
                p0 = (d0 << 11) | f.d0; p1 = (d1 << 11) | f.d1;
                if (first == 0 && threadIdx.x == 0) { st->d2[0] = p0; st->d2[1] = p1; st->r2[0] = f.r0; st->r2[1] = f.r1; }
                clear_hist(2 * NBF3);
            }
            HistOp op{sh.hist, which == 2 ? HIST_X2 : HIST_X3, p0, p1, p0 != p1};
            K::visit(bk, job, first, count, false, ksm, op);
            if (which == 2) flush_bins(histf2(j), 2 * NBF); else flush_bins(histf3(j), 2 * NBF3);
        });
    }

    // ---- A / XA: threshold from the last histogram; apply; counts -----------------------------------------------------------
    __device__ __noinline__ void phase_apply(bool exact) {
        for_each_segment(false, [&](int j, int first, int count) {
            const WtFJob& job = sh.job;
            WtFState* st = state(j);
            float thr;
            if (job.lerp_mode == WT_LERP_ABSOLUTE) {
                if (exact) return;
                thr = job.abs_threshold;
            } else {
                if (exact != (__ldcg(&st->failed) != 0)) return;
                unsigned int key0, key1;
                if (!exact) {
                    const unsigned int lo = __ldcg(&st->lo_key), w = __ldcg(&st->width);
                    const unsigned long long below = __ldcg(&st->below);
                    int s1, take2, s2;
                    tail_shifts(w, s1, take2, s2);
                    unsigned int q0, q1;
                    if (s2 == 0) {
                        const unsigned int d0 = __ldcg(&st->d[0]), d1 = __ldcg(&st->d[1]);
                        const unsigned int* g2 = histf2(j);
                        const Found f = scan_find_pair<T, NBF>(g2, d0 != d1 ? g2 + NBF : g2, 0ull, 0ull, __ldcg(&st->r[0]), __ldcg(&st->r[1]), sh);
                        q0 = (d0 << take2) | f.d0; q1 = (d1 << take2) | f.d1;
                    } else {
                        const unsigned int d0 = __ldcg(&st->d2[0]), d1 = __ldcg(&st->d2[1]);
                        const unsigned int* g3 = histf3(j);
                        const Found f = scan_find_pair<T, NBF3>(g3, d0 != d1 ? g3 + NBF3 : g3, 0ull, 0ull, __ldcg(&st->r2[0]), __ldcg(&st->r2[1]), sh);
                        q0 = (d0 << s2) | f.d0; q1 = (d1 << s2) | f.d1;
                    }
                    unsigned long long e0, e1;
                    explicit_ranks(job, e0, e1);
                    const unsigned long long ez = lo == 1u ? below : 0ull;         // exact zeros among the explicit keys
                    key0 = (job.k0 < job.zeros || e0 < ez) ? 0u : lo + q0;         // a rank inside the zeros is 0
                    key1 = (job.k1 < job.zeros || e1 < ez) ? 0u : lo + q1;
                } else {
                    const unsigned int n0 = __ldcg(&st->d2[0]), n1 = __ldcg(&st->d2[1]);
                    const unsigned long long z = (unsigned long long)job.zeros;
                    const unsigned int* g3 = histf3(j);
                    const Found f = scan_find_pair<T, NBF3>(g3, n0 != n1 ? g3 + NBF3 : g3, n0 == 0u ? z : 0ull, n1 == 0u ? z : 0ull,
                                                            __ldcg(&st->r2[0]), __ldcg(&st->r2[1]), sh);
                    key0 = (n0 << 8) | f.d0; key1 = (n1 << 8) | f.d1;
                }
                thr = lerp_threshold(__uint_as_float(key0), __uint_as_float(key1), job.gamma, job.lerp_mode);
                if (first == 0 && threadIdx.x == 0) {
                    st->a = __uint_as_float(key0); st->b = __uint_as_float(key1); st->t = thr;
                    a.thresholds[job.layer] = thr;
                }
            }
            unsigned int zc = 0, nzb = 0;
            K::apply(bk, job, first, count, thr, ksm, zc, nzb);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                zc += __shfl_down_sync(0xffffffffu, zc, d);
                nzb += __shfl_down_sync(0xffffffffu, nzb, d);
            }
            if ((threadIdx.x & 31) == 0) {
                if (zc) atomicAdd(&st->zc, (unsigned long long)zc);
                if (nzb) atomicAdd(&st->nzb, (unsigned long long)nzb);
            }
            __syncthreads();
            if (threadIdx.x == 0) {   // the CTA that completes the job's last item writes its counts
                __threadfence();
                const unsigned int prev = atomicAdd(&st->done, (unsigned int)count);
                if (prev + (unsigned int)count == (unsigned int)job.items) {
                    __threadfence();
                    const long long numel = K::numel(job);
                    const long long z = (long long)__ldcg(&st->zc);
                    long long* c = a.counts + (size_t)job.layer * 4;
                    c[0] = numel; c[1] = numel - z; c[2] = z; c[3] = (long long)__ldcg(&st->nzb);
                }
            }
        });
    }

    __device__ __forceinline__ void stamp(int i) const {
        if (a.stamps && threadIdx.x == 0) a.stamps[(size_t)blockIdx.x * WT_F_STAMP_SLOTS + i] = now_ns();
    }

    __device__ void run() {
        stamp(0);
        if (a.n_jobs <= WT_F_JIDX) {
            for (int i = threadIdx.x; i < a.n_jobs; i += T) {
                sh.jidx[0][i] = __ldg(&a.jobs[i].item_begin);
                sh.jidx[1][i] = __ldg(&a.jobs[i].s_begin);
            }
            __syncthreads();
        }
        {   // zero the scratch block of the next run (16-byte stores, grid-strided)
            int4* z = reinterpret_cast<int4*>(a.scratch_next);
            const unsigned long long n16 = a.scratch_bytes >> 4;
            for (unsigned long long i = (unsigned long long)blockIdx.x * T + threadIdx.x; i < n16; i += (unsigned long long)gridDim.x * T)
                z[i] = make_int4(0, 0, 0, 0);
        }
        if (a.n_select > 0) {
            cooperative_groups::grid_group grid = cooperative_groups::this_grid();
            phase_hist_top(true);
            stamp(1);
            grid.sync();
            phase_compact();
            stamp(2);
            grid.sync();
            phase_tail(1);
            stamp(3);
            grid.sync();
            if (__ldcg(&header()->wide)) {   // grid-uniform: set during C
                phase_tail(2);
                grid.sync();
            }
            stamp(4);
            phase_apply(false);
            stamp(5);
            if (__ldcg(&header()->fail_count)) {   // grid-uniform: final since the barrier after T1
                phase_hist_top(false);
                grid.sync();
                phase_exact(2);
                grid.sync();
                phase_exact(3);
                grid.sync();
                phase_apply(true);
                stamp(6);
            }
        } else {
            phase_apply(false);
            stamp(5);
        }
    }
};

template <class K>
__global__ void __launch_bounds__(K::THREADS, K::MIN_CTAS) wt_fused_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ WtFArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    using Sh = Shared<K::THREADS>;
    Sh& sh = *reinterpret_cast<Sh*>(smem_raw);
    unsigned char* ksm = smem_raw + ((sizeof(Sh) + 127) / 128) * 128;
    K::init(ksm);
    Cta<K> cta(bk, a, sh, ksm);
    cta.run();
}

template <class K>
constexpr size_t smem_bytes() { return ((sizeof(Shared<K::THREADS>) + 127) / 128) * 128 + K::SMEM_BYTES; }

// grid = what fits the device minus `reserve`, at most one CTA per item; cooperative when the class runs a select
template <class K>
int launch(const WtBank& bank, const WtFArgs& args, int reserve, double share, cudaStream_t stream, int* grid_out) {
    static int blocks_per_sm[64] = {0};  // per device; racing threads compute the same value
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    WT_CHECK(di.device >= 0 && di.device < 64, WT_E_UNSUPPORTED, "device ordinal %d out of range", di.device);
    WT_CHECK(di.coop, WT_E_UNSUPPORTED, "device does not support cooperative launch");
    const size_t smem = smem_bytes<K>();
    if (blocks_per_sm[di.device] == 0) {
        WT_CUDA(cudaFuncSetAttribute(wt_fused_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int nb = 0;
        WT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wt_fused_kernel<K>, K::THREADS, smem));
        WT_CHECK(nb > 0, WT_E_CUDA, "fused kernel does not fit on an SM");
        blocks_per_sm[di.device] = nb;
    }
    const int cap = blocks_per_sm[di.device] * di.sm_count;
    int grid = cap - reserve;
    const int fair = (int)ceil(share * cap);
    if (grid > fair) grid = fair;
    if (grid > args.total_items) grid = args.total_items;
    if (grid < 1) grid = 1;
    if (grid_out) *grid_out = grid;
    if (args.n_select > 0) {
        WtBank b = bank;
        WtFArgs a = args;
        void* params[] = {&b, &a};
        WT_CUDA(cudaLaunchCooperativeKernel((const void*)wt_fused_kernel<K>, dim3(grid), dim3(K::THREADS), params, smem, stream));
    } else {
        wt_fused_kernel<K><<<grid, K::THREADS, smem, stream>>>(bank, args);
        WT_CUDA(cudaGetLastError());
    }
    return WT_OK;
}

}  // namespace wtf
