// wt_fused.cuh -- the fused prune kernel: ONE persistent launch per plane class does the whole
//   wavedec2 -> np.percentile(|c|, p) -> np.where(|c| < t, 0, c) -> waverec2 -> crop -> zero count
// of /root/reference/ResNet/dwt_pruning.py:35-95 (and the absolute-threshold flow of
// dwt_pruning_NoEntropy.py:12-62) for every tensor of the class, WITHOUT ever writing a coefficient to HBM: the
// coefficients of a small plane live in registers and are recomputed from the weights in every pass that needs them
// (hard part 2).  Algorithmic traffic stays what it is -- read the weight, write the pruned weight.
//
// Dataflow instead of grid barriers.  The work of a launch is a static queue of tickets; a CTA draws the next ticket
// with one atomicAdd (prefetched under the previous item) and every job (tensor) walks through its stages on its own:
//   S   12-bit histogram (key bits 30..19 of |c|) of a 1/2^s SAMPLE of the job's keys (whole for jobs < 32 Ki keys)
//       -> the CTA that finishes the job's last S item brackets both ranks in consecutive bins (5 sigma of the
//          binomial rank error) and publishes the bracket
//   C   the one full pass: count the keys below the bracket, compact the keys inside it into the job's candidate
//       list (warp-private staging) and histogram the top 11 bits of (key - lo) of those candidates
//       -> the last CTA verifies the bracket with the EXACT counts (a miss hands the job to the exact path X below, so
//          the result never depends on the sample) and locates both ranks in that histogram
//   F2  next <= 11 bits over the candidate list (a few slices per job, L2-resident)  -> usually final
//   F3  last <= 9 bits (only for brackets wider than 8 bins: a rank next to the massive tie at zero)
//       -> sorted[k0], sorted[k1] -> np.percentile's lerp (legacy_1_22 / nep50_2x) -> threshold
//   A   apply: recompute the coefficients, zero |c| < t, synthesise, crop, store, count zeros
// "->" is done by whichever CTA happens to complete the stage of that job last (a per-job counter), which then raises
// a per-job flag; an item whose job is not ready yet spins on that flag.  Tickets are drawn in queue order and a stage
// only ever waits for tickets that were drawn earlier, so there is no deadlock whatever the number of resident CTAs,
// and nobody waits for anybody else's job: no grid barrier, no cooperative launch, no elected-CTA tail.  The queue is a
// software pipeline over the jobs (C(j), F2(j-1), F3(j-2), A(j-3), smallest job first), so the latency of the
// stage hand-overs of one job hides under the streaming passes of its neighbours and A(j) finds the weights C(j) read
// a moment ago still in L2.
//
// X1..X3 + XA: exact MSD radix select (12 + 11 + 8 bits, three recomputing passes, no candidate list) for the jobs
// whose bracket missed or whose list overflowed; drawn from a second ticket counter after the main queue, by the same
// CTAs, only if some job failed.
//
// The per-run scratch (ticket counters, per-job counters, flags and histograms) must be zero at launch; every launch
// zeroes the scratch block of the NEXT run (two blocks alternate), so a step enqueues no memset.
#pragma once
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
    int ticket;       // next ticket of this CTA (prefetched)
    int flag;         // broadcast slot (last finisher, fallback decision)
    WtFJob job;       // the job of the current context
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

// the histogram passes: S / X1 bin key bits 30..19, X2 bits 18..8 and X3 bits 7..0 of the keys that match the prefix
struct HistOp {
    unsigned int* h;
    int stage;
    unsigned int p0, p1;
    bool two;
    __device__ __forceinline__ void operator()(unsigned int key, bool valid) const {
        if (!valid) return;
        if (stage == WT_ST_S || stage == WT_ST_X1) {
            atomicAdd(&h[key >> 19], 1u);
        } else if (stage == WT_ST_X2) {
            const unsigned int top = key >> 19;
            if (top == p0) atomicAdd(&h[(key >> 8) & 2047u], 1u);
            else if (two && top == p1) atomicAdd(&h[NBF + ((key >> 8) & 2047u)], 1u);
        } else {
            const unsigned int top = key >> 8;
            if (top == p0) atomicAdd(&h[key & 255u], 1u);
            else if (two && top == p1) atomicAdd(&h[NBF3 + (key & 255u)], 1u);
        }
    }
};

// ======================================================================================================================
// The CTA-side state machine.  K (the plane class) provides:
//   THREADS, MIN_CTAS, SMEM_BYTES (class-private shared memory after Shared<THREADS>)
//   init(smem)                                       once per CTA
//   visit(bk, job, item, sampled, smem, op)          op(key, valid) for every explicit key of work item `item` (sampled:
//                                                    of sample item `item`), called by all 32 lanes of a warp together
//   apply(bk, job, item, t, smem, zc, nzb)           out = prune(in) for the item; zc += zeros written, nzb += nonzero inputs
//   numel(job)                                       weight elements of the job
// ======================================================================================================================
template <class K>
struct Cta {
    static constexpr int T = K::THREADS;
    using Sh = Shared<T>;

    const WtBank& bk;
    const WtFArgs& a;
    Sh& sh;
    unsigned char* ksm;  // class-private shared memory
    // current context: every item of (stage, job) since the last flush
    int stage = -1, job_i = -1, pending = 0;
    bool skip = false;   // the job needs nothing from this stage (failed, already final, absolute threshold ...)
    // stage-private accumulators
    unsigned int nbelow = 0, wcount = 0;  // C (wcount is warp-uniform)
    unsigned int zc = 0, nzb = 0;         // A
    // stage parameters read from the job state when the context opens
    unsigned int lo_key = 0, width = 0, pa = 0, pb = 0;
    int sh_hi = 0, sh_lo = 0;
    bool two = false;
    float thr = 0.0f;

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
    __device__ __forceinline__ void wait_flag(const unsigned int* f) {
        if (threadIdx.x == 0)
            while (ld_acquire(f) == 0u) __nanosleep(40);
        __syncthreads();
    }
    __device__ __forceinline__ void warp_flush(unsigned int& wcount) {
        const int lane = threadIdx.x & 31;
        const unsigned int* wbuf = sh.stage + (threadIdx.x >> 5) * WT_F_WSTAGE;
        unsigned int gbase = 0;
        if (lane == 0) gbase = atomicAdd(&state(job_i)->s_count, wcount);
        gbase = __shfl_sync(0xffffffffu, gbase, 0);
        __syncwarp();
        unsigned int* cand = sh.job.cand;
        const unsigned int cap = sh.job.cand_cap;
        for (unsigned int i = lane; i < wcount; i += 32)
            if (gbase + i < cap) cand[gbase + i] = wbuf[i];
        __syncwarp();
        wcount = 0;
    }

    // ---- failure: the job leaves the sampled pipeline; every later stage skips it and X1..XA run it after the queue ----
    __device__ void fail_job(int j) {  // whole CTA
        unsigned int* gh = hist1(j);
        for (int b = threadIdx.x; b < NB1 + NCOARSE; b += T) gh[b] = 0u;  // held the sample: X1 starts from zero
        __syncthreads();
        if (threadIdx.x == 0) {
            WtFState* st = state(j);
            st->failed = 1;
            __threadfence();
            atomicAdd(&header()->fail_count, 1u);
            st_release(&st->ready[WT_RD_S], 1u);
            st_release(&st->ready[WT_RD_C], 1u);
            st_release(&st->ready[WT_RD_F2], 1u);
            st_release(&st->ready[WT_RD_T], 1u);
            __threadfence();
            atomicAdd(&header()->verified, 1u);
        }
    }

    // both order statistics as key patterns -> threshold -> published (thread 0 stores); v0, v1 = key - lo_key
    __device__ void publish_threshold(int j, const WtFJob& job, unsigned int key0, unsigned int key1, unsigned int ready_slot) {
        const float t = lerp_threshold(__uint_as_float(key0), __uint_as_float(key1), job.gamma, job.lerp_mode);
        if (threadIdx.x == 0) {
            WtFState* st = state(j);
            st->a = __uint_as_float(key0);
            st->b = __uint_as_float(key1);
            st->t = t;
            a.thresholds[job.layer] = t;
            st_release(&st->ready[ready_slot], 1u);
        }
    }
    __device__ void finish_sampled(int j, const WtFJob& job, unsigned int v0, unsigned int v1) {
        WtFState* st = state(j);
        const unsigned int lo = __ldcg(&st->lo_key);
        unsigned long long e0, e1;
        explicit_ranks(job, e0, e1);
        const unsigned long long ez = (lo == 1u) ? __ldcg(&st->below) : 0ull;  // exact zeros among the explicit keys
        const unsigned int key0 = (job.k0 < job.zeros || e0 < ez) ? 0u : lo + v0;  // a rank inside the zeros is 0
        const unsigned int key1 = (job.k1 < job.zeros || e1 < ez) ? 0u : lo + v1;
        publish_threshold(j, job, key0, key1, WT_RD_T);
    }

    // ---- what the CTA that completes stage `s` of job j does ("->" in the header comment) ------------------------------
    __device__ void finisher(int s, int j) {
        const WtFJob& job = sh.job;
        WtFState* st = state(j);
        if (s == WT_ST_S) {
            unsigned long long e0, e1;
            explicit_ranks(job, e0, e1);
            const Found f = scan_find_two_level<T>(coarse(j), hist1(j), 0ull, e0, e1, sh, job.n);
            if (f.d1 < f.d0 || job.cand == nullptr) { fail_job(j); return; }
            // a bracket that starts at bin 0 leaves the exact zeros out of the list: they are counted as "below" and a
            // rank inside them is 0 without looking (the massive tie of this domain: pruned weights, vanishing details)
            const unsigned int lo = (f.d0 << 19) | (f.d0 == 0u ? 1u : 0u);
            const unsigned int w = ((f.d1 + 1u) << 19) - lo;
            const int bits = 32 - __clz(w - 1u);
            const int s1 = bits - 11, take2 = s1 < 11 ? s1 : 11;
            if (threadIdx.x == 0) {
                st->lo_key = lo; st->width = w; st->sh1 = s1; st->sh2 = s1 - take2;
                st_release(&st->ready[WT_RD_S], 1u);
            }
        } else if (s == WT_ST_C) {
            if (__ldcg(&st->failed)) return;
            const unsigned int n_c = __ldcg(&st->s_count);
            const unsigned long long below = __ldcg(&st->below);
            unsigned long long e0, e1;
            explicit_ranks(job, e0, e1);
            const bool zlo = __ldcg(&st->lo_key) == 1u;  // then `below` counts the exact zeros
            const bool in0 = (zlo && e0 < below) || (below <= e0 && e0 < below + (unsigned long long)n_c);
            const bool in1 = (zlo && e1 < below) || (below <= e1 && e1 < below + (unsigned long long)n_c);
            if (!(n_c <= job.cand_cap && in0 && in1)) { fail_job(j); return; }
            const unsigned int* h = histf1(j);
            const Found f = scan_find_pair<T, NBF>(h, h, 0ull, 0ull, e0 < below ? 0ull : e0 - below, e1 < below ? 0ull : e1 - below, sh);
            if (threadIdx.x == 0) {
                st->d[0] = f.d0; st->d[1] = f.d1; st->r[0] = f.r0; st->r[1] = f.r1;
                st_release(&st->ready[WT_RD_C], 1u);
                __threadfence();
                atomicAdd(&header()->verified, 1u);
            }
        } else if (s == WT_ST_F2) {
            if (__ldcg(&st->failed)) return;
            const unsigned int d0 = __ldcg(&st->d[0]), d1 = __ldcg(&st->d[1]);
            const int s1 = __ldcg(&st->sh1), s2 = __ldcg(&st->sh2);
            const unsigned int* h = histf2(j);
            const Found f = scan_find_pair<T, NBF>(h, d0 != d1 ? h + NBF : h, 0ull, 0ull, __ldcg(&st->r[0]), __ldcg(&st->r[1]), sh);
            const unsigned int q0 = (d0 << (s1 - s2)) | f.d0, q1 = (d1 << (s1 - s2)) | f.d1;
            if (s2 == 0) {
                finish_sampled(j, job, q0, q1);
            } else if (threadIdx.x == 0) {
                st->d2[0] = q0; st->d2[1] = q1; st->r2[0] = f.r0; st->r2[1] = f.r1;
                st_release(&st->ready[WT_RD_F2], 1u);
            }
        } else if (s == WT_ST_F3) {
            if (__ldcg(&st->failed) || __ldcg(&st->sh2) == 0) return;
            const unsigned int q0 = __ldcg(&st->d2[0]), q1 = __ldcg(&st->d2[1]);
            const int s2 = __ldcg(&st->sh2);
            const unsigned int* h = histf3(j);
            const Found f = scan_find_pair<T, NBF3>(h, q0 != q1 ? h + NBF3 : h, 0ull, 0ull, __ldcg(&st->r2[0]), __ldcg(&st->r2[1]), sh);
            finish_sampled(j, job, (q0 << s2) | f.d0, (q1 << s2) | f.d1);
        } else if (s == WT_ST_X1) {
            const Found f = scan_find_two_level<T>(coarse(j), hist1(j), (unsigned long long)job.zeros, (unsigned long long)job.k0,
                                                   (unsigned long long)job.k1, sh, 0);
            if (threadIdx.x == 0) {
                st->d[0] = f.d0; st->d[1] = f.d1; st->r[0] = f.r0; st->r[1] = f.r1;
                st_release(&st->ready[WT_RD_X1], 1u);
            }
        } else if (s == WT_ST_X2) {
            const unsigned int p0 = __ldcg(&st->d[0]), p1 = __ldcg(&st->d[1]);
            const unsigned long long z = (unsigned long long)job.zeros;
            const unsigned int* h = histf2(j);
            const Found f = scan_find_pair<T, NBF>(h, p0 != p1 ? h + NBF : h, p0 == 0u ? z : 0ull, p1 == 0u ? z : 0ull,
                                                   __ldcg(&st->r[0]), __ldcg(&st->r[1]), sh);
            if (threadIdx.x == 0) {
                st->d2[0] = (p0 << 11) | f.d0; st->d2[1] = (p1 << 11) | f.d1; st->r2[0] = f.r0; st->r2[1] = f.r1;
                st_release(&st->ready[WT_RD_X2], 1u);
            }
        } else if (s == WT_ST_X3) {
            const unsigned int n0 = __ldcg(&st->d2[0]), n1 = __ldcg(&st->d2[1]);
            const unsigned long long z = (unsigned long long)job.zeros;
            const unsigned int* h = histf3(j);
            const Found f = scan_find_pair<T, NBF3>(h, n0 != n1 ? h + NBF3 : h, n0 == 0u ? z : 0ull, n1 == 0u ? z : 0ull,
                                                    __ldcg(&st->r2[0]), __ldcg(&st->r2[1]), sh);
            publish_threshold(j, job, (n0 << 8) | f.d0, (n1 << 8) | f.d1, WT_RD_XT);
        } else {  // A / XA: the job's counts are complete
            if (s == WT_ST_A && job.lerp_mode != WT_LERP_ABSOLUTE && __ldcg(&st->failed)) return;
            if (threadIdx.x == 0) {
                const long long numel = K::numel(job);
                const long long z = (long long)__ldcg(&st->zc);
                long long* c = a.counts + (size_t)job.layer * 4;
                c[0] = numel; c[1] = numel - z; c[2] = z; c[3] = (long long)__ldcg(&st->nzb);
            }
        }
    }

    // number of items stage s has for the job
    __device__ __forceinline__ int stage_items(int s, const WtFJob& job) const {
        if (s == WT_ST_S) return job.s_chunks;
        if (s == WT_ST_F2 || s == WT_ST_F3) return job.f_chunks;
        return job.chunks;
    }

    // ---- closes the current context: partial results to the job's global state, then the completion count ------------
    __device__ __noinline__ void flush() {
        if (stage < 0) return;
        const int s = stage, j = job_i;
        WtFState* st = state(j);
        if (!skip) {
            if (s == WT_ST_S || s == WT_ST_X1) {
                flush_hist1(j);
            } else if (s == WT_ST_C) {
                if (wcount) warp_flush(wcount);
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) nbelow += __shfl_down_sync(0xffffffffu, nbelow, d);
                if ((threadIdx.x & 31) == 0 && nbelow) atomicAdd(&st->below, (unsigned long long)nbelow);
                nbelow = 0;
                flush_bins(histf1(j), NBF);
            } else if (s == WT_ST_F2 || s == WT_ST_X2) {
                flush_bins(histf2(j), 2 * NBF);
            } else if (s == WT_ST_F3 || s == WT_ST_X3) {
                flush_bins(histf3(j), 2 * NBF3);
            } else {  // A / XA
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) {
                    zc += __shfl_down_sync(0xffffffffu, zc, d);
                    nzb += __shfl_down_sync(0xffffffffu, nzb, d);
                }
                if ((threadIdx.x & 31) == 0) {
                    if (zc) atomicAdd(&st->zc, (unsigned long long)zc);
                    if (nzb) atomicAdd(&st->nzb, (unsigned long long)nzb);
                }
                zc = nzb = 0;
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned int prev = atomicAdd(&st->done[s], (unsigned int)pending);
            const int last = (prev + (unsigned int)pending == (unsigned int)stage_items(s, sh.job));
            if (last) __threadfence();
            sh.flag = last;
        }
        __syncthreads();
        const int last = sh.flag;
        __syncthreads();
        if (last) finisher(s, j);
        stage = -1;
        pending = 0;
    }

    // ---- opens the context (stage s, job j): dependencies, stage parameters, shared histogram ----------------------------
    __device__ void open(int s, int j) {
        stage = s; job_i = j; pending = 0; skip = false;
        __syncthreads();
        {   // the job descriptor into shared memory (int4 copies)
            const int4* src = reinterpret_cast<const int4*>(a.jobs + j);
            int4* dst = reinterpret_cast<int4*>(&sh.job);
            for (int i = threadIdx.x; i < (int)(sizeof(WtFJob) / 16); i += T) dst[i] = __ldg(&src[i]);
        }
        __syncthreads();
        const WtFJob& job = sh.job;
        WtFState* st = state(j);
        const bool absolute = job.lerp_mode == WT_LERP_ABSOLUTE;
        switch (s) {
            case WT_ST_S:
                clear_hist(NB1);
                break;
            case WT_ST_C:
                wait_flag(&st->ready[WT_RD_S]);
                if (__ldcg(&st->failed)) { skip = true; break; }
                lo_key = __ldcg(&st->lo_key); width = __ldcg(&st->width); sh_hi = __ldcg(&st->sh1);
                nbelow = 0; wcount = 0;
                clear_hist(NBF);
                break;
            case WT_ST_F2:
                wait_flag(&st->ready[WT_RD_C]);
                if (__ldcg(&st->failed)) { skip = true; break; }
                lo_key = __ldcg(&st->lo_key); sh_hi = __ldcg(&st->sh1); sh_lo = __ldcg(&st->sh2);
                pa = __ldcg(&st->d[0]); pb = __ldcg(&st->d[1]); two = pa != pb;
                clear_hist(2 * NBF);
                break;
            case WT_ST_F3:
                wait_flag(&st->ready[WT_RD_C]);   // sh2 is known since S; F2's digits are needed only if F3 runs
                if (__ldcg(&st->failed) || __ldcg(&st->sh2) == 0) { skip = true; break; }
                wait_flag(&st->ready[WT_RD_F2]);
                lo_key = __ldcg(&st->lo_key); sh_lo = __ldcg(&st->sh2);
                pa = __ldcg(&st->d2[0]); pb = __ldcg(&st->d2[1]); two = pa != pb;
                clear_hist(2 * NBF3);
                break;
            case WT_ST_A:
                if (absolute) { thr = job.abs_threshold; zc = nzb = 0; break; }
                wait_flag(&st->ready[WT_RD_T]);
                if (__ldcg(&st->failed)) { skip = true; break; }
                thr = __ldcg(&st->t); zc = nzb = 0;
                break;
            case WT_ST_X1:
                clear_hist(NB1);
                break;
            case WT_ST_X2:
                wait_flag(&st->ready[WT_RD_X1]);
                pa = __ldcg(&st->d[0]); pb = __ldcg(&st->d[1]); two = pa != pb;
                clear_hist(2 * NBF);
                break;
            case WT_ST_X3:
                wait_flag(&st->ready[WT_RD_X2]);
                pa = __ldcg(&st->d2[0]); pb = __ldcg(&st->d2[1]); two = pa != pb;
                clear_hist(2 * NBF3);
                break;
            default:  // XA
                wait_flag(&st->ready[WT_RD_XT]);
                thr = __ldcg(&st->t); zc = nzb = 0;
                break;
        }
    }

    // ---- one item ----------------------------------------------------------------------------------------------------
    __device__ __forceinline__ void run_item(int s, int j, int idx) {
        if (s != stage || j != job_i) {
            flush();
            open(s, j);
        }
        ++pending;
        if (skip) return;
        const WtFJob& job = sh.job;
        unsigned int* h = sh.hist;
        switch (s) {
            case WT_ST_S:
            case WT_ST_X1:
            case WT_ST_X2:
            case WT_ST_X3: {   // one instantiation of the class's visit for the four histogram passes
                HistOp op{h, s, pa, pb, two};
                K::visit(bk, job, idx, s == WT_ST_S, ksm, op);
            } break;
            case WT_ST_C: {
                const int lane = threadIdx.x & 31;
                unsigned int* wbuf = sh.stage + (threadIdx.x >> 5) * WT_F_WSTAGE;
                const unsigned int lo = lo_key, w = width;
                const int s1 = sh_hi;
                const unsigned int lt = (1u << lane) - 1u;
                unsigned int nb = nbelow, wc = wcount;   // hot accumulators in registers for the pass
                auto op = [&, h, wbuf, lo, w, s1, lt](unsigned int key, bool valid) {
                    nb += (valid && key < lo);
                    const unsigned int v = key - lo;
                    const bool hit = valid && v < w;
                    const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
                    if (ballot) {
                        if (hit) {
                            wbuf[wc + __popc(ballot & lt)] = key;
                            atomicAdd(&h[v >> s1], 1u);
                        }
                        wc += __popc(ballot);
                        if (wc > WT_F_WFLUSH) warp_flush(wc);
                    }
                };
                K::visit(bk, job, idx, false, ksm, op);
                nbelow = nb;
                wcount = wc;
            } break;
            case WT_ST_F2:
            case WT_ST_F3: {
                WtFState* st = state(j);
                const unsigned int n_c = __ldcg(&st->s_count);
                const unsigned int lo = cand_boundary(n_c, idx, job.f_chunks), hi = cand_boundary(n_c, idx + 1, job.f_chunks);
                const unsigned int lk = lo_key, p0 = pa, p1 = pb;
                const bool tw = two;
                if (s == WT_ST_F2) {
                    const int s1 = sh_hi, s2 = sh_lo;
                    const unsigned int mask = (1u << (s1 - s2)) - 1u;
                    cand_for_each<T>(job.cand + lo, hi - lo, [=](unsigned int k) {
                        const unsigned int v = k - lk, top = v >> s1;
                        if (top == p0) atomicAdd(&h[(v >> s2) & mask], 1u);
                        else if (tw && top == p1) atomicAdd(&h[NBF + ((v >> s2) & mask)], 1u);
                    });
                } else {
                    const int s2 = sh_lo;
                    const unsigned int mask = (1u << s2) - 1u;
                    cand_for_each<T>(job.cand + lo, hi - lo, [=](unsigned int k) {
                        const unsigned int v = k - lk, top = v >> s2;
                        if (top == p0) atomicAdd(&h[v & mask], 1u);
                        else if (tw && top == p1) atomicAdd(&h[NBF3 + (v & mask)], 1u);
                    });
                }
            } break;
            default: {  // A / XA
                unsigned int z = zc, nb = nzb;
                K::apply(bk, job, idx, thr, ksm, z, nb);
                zc = z;
                nzb = nb;
            } break;
        }
    }

    // ticket -> (stage, job, index) through the segment table; `seg` caches the last hit (tickets only grow)
    __device__ __forceinline__ void decode(int t, int& seg, int& s, int& j, int& idx) const {
        int first = __ldg(&a.segs[seg].first), count = __ldg(&a.segs[seg].count);
        if (t >= first + count) {
            int lo = seg + 1, hi = a.n_segs - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (__ldg(&a.segs[mid].first) <= t) lo = mid; else hi = mid - 1;
            }
            seg = lo;
            first = __ldg(&a.segs[seg].first);
        }
        s = __ldg(&a.segs[seg].stage);
        j = __ldg(&a.segs[seg].job);
        idx = t - first;
    }

    // fallback ticket -> (stage, job, index): stage-major over the failed jobs (rare path, linear walk)
    __device__ bool decode_fallback(int t, int& s, int& j, int& idx) const {
        for (int st = WT_ST_X1; st <= WT_ST_XA; ++st)
            for (int q = 0; q < a.n_jobs; ++q) {
                if (!__ldcg(&state(q)->failed)) continue;
                const int n = __ldg(&a.jobs[q].chunks);
                if (t < n) { s = st; j = q; idx = t; return true; }
                t -= n;
            }
        return false;
    }

    __device__ void run() {
        // zero the scratch block of the next run (16-byte stores, grid-strided)
        {
            int4* z = reinterpret_cast<int4*>(a.scratch_next);
            const unsigned long long n16 = a.scratch_bytes >> 4;
            for (unsigned long long i = (unsigned long long)blockIdx.x * T + threadIdx.x; i < n16; i += (unsigned long long)gridDim.x * T)
                z[i] = make_int4(0, 0, 0, 0);
        }
        WtFHeader* hd = header();
        if (threadIdx.x == 0) sh.ticket = atomicAdd(&hd->ticket, 1);
        __syncthreads();
        int seg = 0;
        bool fallback = false;
        while (true) {
            const int t = sh.ticket;
            __syncthreads();
            int next = 0, s, j, idx;
            if (!fallback) {
                if (t >= a.n_tickets) {
                    flush();
                    if (a.n_select == 0) return;
                    // every select job has been verified (or has failed) before the jobs that failed are counted
                    if (threadIdx.x == 0) {
                        while (ld_acquire(&hd->verified) < (unsigned int)a.n_select) __nanosleep(100);
                        sh.flag = (int)ld_acquire(&hd->fail_count);
                    }
                    __syncthreads();
                    const int failed = sh.flag;
                    __syncthreads();
                    if (!failed) return;
                    fallback = true;   // X1..XA of the failed jobs, from the second ticket counter
                    if (threadIdx.x == 0) sh.ticket = atomicAdd(&hd->ticket2, 1);
                    __syncthreads();
                    continue;
                }
                if (threadIdx.x == 0) next = atomicAdd(&hd->ticket, 1);   // in flight while this item runs
                decode(t, seg, s, j, idx);
            } else {
                if (!decode_fallback(t, s, j, idx)) break;
                if (threadIdx.x == 0) next = atomicAdd(&hd->ticket2, 1);
            }
            run_item(s, j, idx);
            if (threadIdx.x == 0) sh.ticket = next;
            __syncthreads();
        }
        flush();
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

// launch helper: grid = min(tickets, resident CTAs)
template <class K>
int launch(const WtBank& bank, const WtFArgs& args, int max_grid, cudaStream_t stream) {
    static int blocks_per_sm[64] = {0};  // per device; racing threads compute the same value
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    WT_CHECK(di.device >= 0 && di.device < 64, WT_E_UNSUPPORTED, "device ordinal %d out of range", di.device);
    const size_t smem = smem_bytes<K>();
    if (blocks_per_sm[di.device] == 0) {
        WT_CUDA(cudaFuncSetAttribute(wt_fused_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int nb = 0;
        WT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wt_fused_kernel<K>, K::THREADS, smem));
        WT_CHECK(nb > 0, WT_E_CUDA, "fused kernel does not fit on an SM");
        blocks_per_sm[di.device] = nb;
    }
    int grid = blocks_per_sm[di.device] * di.sm_count;
    if (max_grid > 0 && grid > max_grid) grid = max_grid;
    if (grid > args.n_tickets) grid = args.n_tickets;
    if (grid < 1) grid = 1;
    wt_fused_kernel<K><<<grid, K::THREADS, smem, stream>>>(bank, args);
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

}  // namespace wtf
