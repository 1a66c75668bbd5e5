// wt_fused_kinds.cuh -- the plane classes the fused prune kernel (wt_fused.cuh) is instantiated for.
//
//   FlatKind            level-0 jobs of any shape: wavedec2(level=0) returns [data], so the "DWT prune" of
//                       /root/reference/ResNet/dwt_pruning.py:64-77 degenerates to select-and-zero on the raw weights
//                       (SURVEY F5); also <2-D tensors (dwt_pruning.py:58-62).  Keys = |w|.
//   SmallKind<H,W,F,Z>  conv-kernel planes 1x1 / 3x3 / 5x5 / 7x7 at levels 1..6, periodization: one thread per plane, the
//                       whole multi-level transform in registers (wt_small_core.cuh), planes staged through shared memory
//                       with 1-D TMA bulk copies (cp.async.bulk + mbarrier, double-buffered over a segment's tiles).  Keys = |c| of
//                       the coefficients, recomputed from the weights wherever they are needed; never stored.
#pragma once
#include "wt_fused.cuh"
#include "wt_small_core.cuh"

namespace wtf {

// ======================================================================================================================
struct FlatKind {
    static constexpr int THREADS = 256;
    static constexpr int MIN_CTAS = 4;
    static constexpr size_t SMEM_BYTES = 0;
    static constexpr int UNROLL = 4;                 // float4 loads in flight per thread
    static constexpr int CHUNK = WT_F_FLAT_CHUNK;    // elements per item: one trip of the vector loop
    static_assert(CHUNK == THREADS * UNROLL * 4, "one item = one trip");

    __device__ static void init(unsigned char*) {}
    __device__ static long long numel(const WtFJob& job) { return job.units; }

    // op(keys[16]) over every element of [x, x + len): all lanes call together, trip counts are CTA-uniform
    template <class Op>
    __device__ __forceinline__ static void span(const float* x, long long len, Op& op) {
        long long done = 0;
        if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
            const long long nvec = len >> 2;
            const float4* x4 = reinterpret_cast<const float4*>(x);
            for (long long v0 = 0; v0 < nvec; v0 += THREADS * UNROLL) {
                const long long v = v0 + threadIdx.x;
                float4 r[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if (v + u * THREADS < nvec) r[u] = __ldcg(&x4[v + u * THREADS]);
                unsigned int key[UNROLL * 4];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    const bool ok = v + u * THREADS < nvec;
                    key[u * 4 + 0] = ok ? abs_key(r[u].x) : WT_F_NOKEY; key[u * 4 + 1] = ok ? abs_key(r[u].y) : WT_F_NOKEY;
                    key[u * 4 + 2] = ok ? abs_key(r[u].z) : WT_F_NOKEY; key[u * 4 + 3] = ok ? abs_key(r[u].w) : WT_F_NOKEY;
                }
                op(key);
            }
            done = nvec << 2;
        }
        for (long long i0 = done; i0 < len; i0 += THREADS) {  // tail, or everything when x is not 16-byte aligned
            unsigned int key[1];
            key[0] = (i0 + threadIdx.x < len) ? abs_key(__ldcg(&x[i0 + threadIdx.x])) : WT_F_NOKEY;
            op(key);
        }
    }

    template <class Op>
    __device__ __forceinline__ static void visit(const WtBank&, const WtFJob& job, int first, int count, bool sampled, unsigned char*,
                                                 Op& op) {
        const int s = job.sample_shift;
        if (!sampled || s == 0) {
            const long long base = (long long)first * CHUNK;
            long long len = job.units - base;
            if (len > (long long)count * CHUNK) len = (long long)count * CHUNK;
            span(job.in + base, len, op);
            return;
        }
        // sample item = 16 tiles of 4096 elements: a window of 4096 >> s elements out of every tile, at an offset that
        // rotates with the tile index
        const long long n_tiles = (job.units + 4095) >> 12;
        const long long tile0 = (long long)first * 16;
        long long tiles = n_tiles - tile0;
        if (tiles > (long long)count * 16) tiles = (long long)count * 16;
        const float* x = job.in + (tile0 << 12);
        const long long len = job.units - (tile0 << 12);
        const int wshift = 12 - s;      // log2(window elements)
        const int vshift = wshift - 2;  // log2(float4s per window)
        const unsigned int wmask = (1u << s) - 1u, vmask = (1u << vshift) - 1u;
        const long long total = tiles << vshift;
        const bool aligned = (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
        for (long long v0 = 0; v0 < total; v0 += THREADS * UNROLL) {
            unsigned int key[UNROLL * 4];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const long long v = v0 + threadIdx.x + u * THREADS;
                float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
                bool ok = false;
                if (v < total) {
                    const long long tile = v >> vshift;
                    const unsigned int win = ((unsigned int)(tile0 + tile) * 5u + 3u) & wmask;
                    const long long off = (tile << 12) + ((long long)win << wshift) + (long long)(((unsigned int)v & vmask) << 2);
                    if (off + 4 <= len) {
                        ok = true;
                        if (aligned) r = __ldcg(reinterpret_cast<const float4*>(x + off));
                        else r = make_float4(__ldcg(x + off), __ldcg(x + off + 1), __ldcg(x + off + 2), __ldcg(x + off + 3));
                    }
                }
                key[u * 4 + 0] = ok ? abs_key(r.x) : WT_F_NOKEY; key[u * 4 + 1] = ok ? abs_key(r.y) : WT_F_NOKEY;
                key[u * 4 + 2] = ok ? abs_key(r.z) : WT_F_NOKEY; key[u * 4 + 3] = ok ? abs_key(r.w) : WT_F_NOKEY;
            }
            op(key);
        }
    }

    __device__ __forceinline__ static float prune_one(float v, float t, unsigned int& zc, unsigned int& nzb) {
        const float o = (fabsf(v) < t) ? 0.0f : v;  // np.where(|c| < t, 0, c), strict (dwt_pruning.py:31)
        zc += (o == 0.0f);
        nzb += (v != 0.0f);
        return o;
    }

    // out = |x| < t ? 0 : x for the items; x may alias out (in-place callers), so no read-only path.  Layer-sharded runs
    // store every value to the same offset of each peer's buffer too, straight from the registers.
    __device__ __forceinline__ static void apply(const WtBank&, const WtFJob& job, int first, int count, float t, unsigned char*,
                                                 unsigned int& zc, unsigned int& nzb) {
        const long long base = (long long)first * CHUNK;
        long long len = job.units - base;
        if (len > (long long)count * CHUNK) len = (long long)count * CHUNK;
        const float* x = job.in + base;
        float* o = job.out + base;
        const int np = job.peers.n;
        long long done = 0;
        if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(o)) & 15u) == 0) {
            const long long nvec = len >> 2;
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float4* o4 = reinterpret_cast<float4*>(o);
            for (long long v = threadIdx.x; v < nvec; v += THREADS * UNROLL) {
                float4 r[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; ++u)
                    if (v + u * THREADS < nvec) r[u] = __ldcg(&x4[v + u * THREADS]);
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    if (v + u * THREADS < nvec) {
                        r[u].x = prune_one(r[u].x, t, zc, nzb); r[u].y = prune_one(r[u].y, t, zc, nzb);
                        r[u].z = prune_one(r[u].z, t, zc, nzb); r[u].w = prune_one(r[u].w, t, zc, nzb);
                        __stcs(&o4[v + u * THREADS], r[u]);
                        for (int q = 0; q < np; ++q)
                            __stcs(reinterpret_cast<float4*>(reinterpret_cast<char*>(&o4[v + u * THREADS]) + job.peers.delta[q]), r[u]);
                    }
                }
            }
            done = nvec << 2;
        }
        for (long long i = done + threadIdx.x; i < len; i += THREADS) {
            const float w = prune_one(x[i], t, zc, nzb);
            o[i] = w;
            for (int q = 0; q < np; ++q) *reinterpret_cast<float*>(reinterpret_cast<char*>(&o[i]) + job.peers.delta[q]) = w;
        }
    }
};

// ======================================================================================================================
template <int H, int W, int F, class Z>
struct SmallKind {
    static constexpr int HW = H * W;
    static constexpr int THREADS = WT_F_SMALL_THREADS(H, W);
    static constexpr int MIN_CTAS = (HW <= 9) ? 3 : 1;
    static constexpr int TILE = THREADS;   // planes per item: one per thread
    static constexpr int NCOEF_MAX = [] {
        int n = 0, h = H, w = W;
        for (int l = 0; l < WT_SMALL_MAX_LEVEL; ++l) { h = (h + 1) / 2; w = (w + 1) / 2; n += 3 * h * w; }
        return n + h * w;
    }();

    // input ring: NST tiles in flight per CTA (the persistent CTA has few warps to hide the HBM latency of a tile behind;
    // the ring does it: bytes in flight per SM = CTAs x (NST - 1) x tile bytes)
    static constexpr int NST = (HW <= 9) ? 3 : 2;
    struct Smem {
        float in[NST][TILE * HW];
        float out[TILE * HW];
        unsigned long long bar[NST];
        unsigned int phase;   // bit b: parity the next wait on bar[b] uses (thread 0 owns it, re-read after a CTA barrier)
    };
    static constexpr size_t SMEM_BYTES = sizeof(Smem);

    __device__ static void init(unsigned char* ksm) {
        Smem& sm = *reinterpret_cast<Smem*>(ksm);
        if (threadIdx.x == 0) {
            for (int b = 0; b < NST; ++b) mbar_init(&sm.bar[b], 1);
            mbar_fence_init();
            sm.phase = 0u;
        }
        __syncthreads();
    }
    __device__ static long long numel(const WtFJob& job) { return job.units * HW; }

    // tile of (sample) item q: the sample takes one tile out of every 2^s, at an offset that rotates with the group index
    __device__ __forceinline__ static long long tile_of(const WtFJob& job, long long n_tiles, long long q, bool sampled) {
        const int s = job.sample_shift;
        if (!sampled || s == 0) return q;
        const long long g0 = q << s;
        long long avail = n_tiles - g0;
        if (avail > (1LL << s)) avail = 1LL << s;
        return g0 + (long long)(((unsigned long long)q * 5ull + 3ull) % (unsigned long long)avail);
    }

    // starts the copy of `tile` into buffer b; returns true when it went out as a TMA bulk copy (CTA-uniform)
    __device__ __forceinline__ static bool issue(const WtFJob& job, Smem& sm, long long tile, int b) {
        const long long p0 = tile * TILE;
        const long long left = job.units - p0;
        const int cnt = left < TILE ? (int)left : TILE;
        const float* g = job.in + p0 * HW;
        const unsigned int bytes = (unsigned int)cnt * HW * 4u;
        const bool tma_ok = ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) && ((bytes & 15u) == 0);
        if (tma_ok) {
            if (threadIdx.x == 0) bulk_load(sm.in[b], g, bytes, &sm.bar[b]);
        } else {
            for (int i = threadIdx.x; i < cnt * HW; i += THREADS) sm.in[b][i] = __ldcg(&g[i]);
        }
        return tma_ok;
    }

    // body(x, valid, cnt, p0) for the tile of every item in [first, first + count), x = this thread's plane (zeros when
    // !valid); the copies of the next NST - 1 tiles are in flight while body runs.  body must contain a CTA barrier after
    // which no thread reads the input ring (x is in registers by then): it frees the buffer for the copy issued next.
    template <class Body>
    __device__ __forceinline__ static void stream_tiles(const WtFJob& job, int first, int count, bool sampled, Smem& sm, Body& body) {
        if (count <= 0) return;
        const long long n_tiles = (job.units + TILE - 1) / TILE;
        unsigned int phase = sm.phase;   // stable: thread 0 updates it only after the last CTA barrier of this function
        unsigned int tma = 0u;           // bit b: the tile in buffer b came by TMA (wait on its mbarrier)
#pragma unroll
        for (int p = 0; p < NST - 1; ++p)
            if (p < count && issue(job, sm, tile_of(job, n_tiles, first + p, sampled), p)) tma |= 1u << p;
        int b = 0, bn = NST - 1;   // buffer of tile i / of tile i + NST - 1
        for (int i = 0; i < count; ++i) {
            if (i + NST - 1 < count) {   // buffer bn held tile i - 1: every thread passed the barrier in its body
                if (issue(job, sm, tile_of(job, n_tiles, first + i + NST - 1, sampled), bn)) tma |= 1u << bn; else tma &= ~(1u << bn);
            }
            if (tma & (1u << b)) {
                mbar_wait(&sm.bar[b], (phase >> b) & 1u);
                phase ^= 1u << b;
            } else {
                __syncthreads();
            }
            const long long p0 = tile_of(job, n_tiles, first + i, sampled) * TILE;
            const long long left = job.units - p0;
            const int cnt = left < TILE ? (int)left : TILE;
            const bool valid = (int)threadIdx.x < cnt;
            float x[HW];
#pragma unroll
            for (int e = 0; e < HW; ++e) x[e] = valid ? sm.in[b][threadIdx.x * HW + e] : 0.0f;
            body(x, valid, cnt, p0);
            b = (b + 1 == NST) ? 0 : b + 1;
            bn = (bn + 1 == NST) ? 0 : bn + 1;
        }
        if (threadIdx.x == 0) sm.phase = phase;
        __syncthreads();
    }

    template <class Op>
    __device__ __forceinline__ static void visit(const WtBank& bk, const WtFJob& job, int first, int count, bool sampled,
                                                 unsigned char* ksm, Op& op) {
        Smem& sm = *reinterpret_cast<Smem*>(ksm);
        const int level = job.level;
        auto body = [&](float (&x)[HW], bool valid, int, long long) {
            __syncthreads();   // every thread holds its plane: the buffer may be refilled
            unsigned int key[NCOEF_MAX];
#pragma unroll
            for (int i = 0; i < NCOEF_MAX; ++i) key[i] = WT_F_NOKEY;   // coefficients the job's level does not have
            auto put = [&](int i, float v) { key[i] = valid ? abs_key(v) : WT_F_NOKEY; };
            wtsmall::SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::analyze(x, level, bk, put);
            op(key);
        };
        stream_tiles(job, first, count, sampled, sm, body);
    }

    __device__ __forceinline__ static void apply(const WtBank& bk, const WtFJob& job, int first, int count, float t, unsigned char* ksm,
                                                 unsigned int& zc, unsigned int& nzb) {
        Smem& sm = *reinterpret_cast<Smem*>(ksm);
        const int level = job.level;
        auto body = [&](float (&x)[HW], bool valid, int cnt, long long p0) {
            if (threadIdx.x == 0) bulk_wait_read();   // the previous tile's bulk store has finished reading sm.out
            __syncthreads();
            float c[NCOEF_MAX];
            auto put = [&](int i, float v) { c[i] = (fabsf(v) < t) ? 0.0f : v; };   // np.where(|c| < t, 0, c), strict
            wtsmall::SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::analyze(x, level, bk, put);
            float y[HW];
            auto get = [&](int i) { return c[i]; };
            wtsmall::SmallPlane<H, W, F, WT_SMALL_MAX_LEVEL, 0, Z>::synthesize(y, level, bk, get);
            if (valid) {
#pragma unroll
                for (int e = 0; e < HW; ++e) {
                    nzb += (x[e] != 0.0f);
                    zc += (y[e] == 0.0f);
                    sm.out[threadIdx.x * HW + e] = y[e];
                }
            }
            float* g = job.out + p0 * HW;
            const unsigned int bytes = (unsigned int)cnt * HW * 4u;
            const bool tma_ok = ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) && ((bytes & 15u) == 0);
            const int np = job.peers.n;
            if (tma_ok) {
                fence_async_smem();
                __syncthreads();
                if (threadIdx.x == 0) {
                    bulk_store(g, sm.out, bytes);
                    // layer-sharded runs: the same tile goes to the same slot of every peer's buffer, as bulk copies
                    // over NVLink issued by this one thread (the SMs do not wait on peer stores)
                    for (int q = 0; q < np; ++q) bulk_store(reinterpret_cast<char*>(g) + job.peers.delta[q], sm.out, bytes);
                    bulk_commit();
                }
            } else {
                __syncthreads();
                for (int i = threadIdx.x; i < cnt * HW; i += THREADS) {
                    const float w = sm.out[i];
                    g[i] = w;
                    for (int q = 0; q < np; ++q) *reinterpret_cast<float*>(reinterpret_cast<char*>(&g[i]) + job.peers.delta[q]) = w;
                }
            }
        };
        stream_tiles(job, first, count, false, sm, body);
        if (threadIdx.x == 0) bulk_wait_all();   // sm.out is free and the stores are performed before the items count as done
        __syncthreads();
    }
};

}  // namespace wtf
