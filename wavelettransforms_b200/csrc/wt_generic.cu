// wt_generic.cu -- generic tiled single-level 2-D DWT / IDWT kernels: any plane size, modes
// 'periodization' and 'symmetric', any of the 17 filter banks.  Replaces pywt.dwt2 / pywt.idwt2
// as iterated by pywt.wavedec2 / waverec2 (/root/reference/ResNet/dwt_pruning.py:67-68,77;
// dwt_pruning_NoEntropy.py:41,49) for the shapes the register-resident small-plane path does not
// cover: large planes (the 4096 x 4096 bank), symmetric mode, odd shapes.
//
// One CTA produces a tile of all four sub-bands (analysis) or of the reconstructed plane
// (synthesis) for `pb` planes: the boundary-extended input region is staged in shared memory
// (the index map handles wrap-around and reflection for any N, including N < F), the first axis
// pass writes its intermediate to shared memory, the second pass writes the result.  Arithmetic and
// accumulation order are those of oracle/dwt_oracle.c (fp32, no FMA).
#include <cstdlib>
#include <cstring>

#include <cuda.h>   // CUtensorMap and the cuTensorMapEncodeTiled prototype only: the entry point is fetched at run time

#include "wt_generic.cuh"

// This file is compiled twice (wt_generic_fma.cu defines WT_GENERIC_FMA and includes it): the faithful arithmetic -- multiply,
// then add, like PyWavelets' C loops without contraction -- and the named fast mode fma_taporder (include/wtprune.h), one
// FMA per tap in the same tap order (oracle twin: orc_set_arithmetic(1)).  Boundary extrapolation (smooth / antireflect
// modes) is not a tap and stays multiply-then-add in both.
#ifdef WT_GENERIC_FMA
#define WT_MAC(acc, tap, v) __fmaf_rn((tap), (v), (acc))
#define WT_GEN_FN(name) name##_fma
#else
#define WT_MAC(acc, tap, v) __fadd_rn((acc), __fmul_rn((tap), (v)))
#define WT_GEN_FN(name) name
#endif

namespace {

constexpr int GEN_THREADS = 256;
constexpr int SMEM_FLOATS = 27000;  // up to 108 KB of dynamic shared memory per CTA (2 CTAs per SM)

// 4-byte asynchronous global -> shared copy (SASS LDGSTS): a thread issues its whole share of a tile
// back to back and waits once, so the staging loop is not bound by one load latency per element
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ---- TMA tensor-map tile loads (cp.async.bulk.tensor, SASS UTMALDG) ------------------------------------------------
// The fixed-geometry kernels stage a whole tile with ONE instruction: a 3-D box {columns, rows, 1 plane} of the strided
// view, written densely into shared memory, out-of-bounds elements filled with zeros by the hardware.  Tiles that
// touch the plane's border then overwrite the out-of-bounds part with the boundary extension (wrap / reflect) from
// plain loads: the map cannot express it, and only ~9 % of the tiles of a 4096 x 4096 plane are border tiles.
__device__ __forceinline__ unsigned int smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect(unsigned long long* bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
__device__ __forceinline__ void tma_load_box(float* smem_dst, const CUtensorMap* map, int col, int row, int plane, unsigned long long* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(map), "r"(col), "r"(row), "r"(plane), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ int pmodi(int a, int m) {
    int r = a % m;
    return r < 0 ? r + m : r;
}

// analysis: index of the boundary-extended signal (interior indices take the first branch)
template <int MODE>
__device__ __forceinline__ int ext_index(int idx, int n) {
    if ((unsigned)idx < (unsigned)n) return idx;
    if (MODE == WT_MODE_PERIODIZATION) {
        int np = n + (n & 1);
        int m = pmodi(idx, np);
        return m == n ? n - 1 : m;
    }
    int m = pmodi(idx, 2 * n);
    return m >= n ? 2 * n - 1 - m : m;
}

struct AnaParams {
    WtView in, aa, da, ad, dd;
    long long batch;
    int h, w, h2, w2, mode;
    int t2r, t2c, tiles_r, tiles_c, pb, ir, ic;
    int ics;  // staged row width: ic + up to 3 columns so that the staged region starts 16-byte aligned, rounded to 4
    int use_tma;  // fixed kernel: the tile comes in through the tensor map
    long long* nz_counts4;  // fixed kernel, level 1: counts4[3] += nonzero inputs, counts4[0] = numel (else nullptr)
};

constexpr int NWARP = GEN_THREADS / 32;

// One CTA: all four sub-bands of a t2r x t2c tile for `pb` planes.  Filter length and mode are template
// constants (taps come straight from the constant bank, loops fully unrolled); warps walk rows, lanes walk
// columns (no integer division anywhere, shared-memory accesses conflict-free: the axis -2 result is stored
// with even and odd columns de-interleaved so that the stride-2 reads of the axis -1 pass are unit-stride).
template <int F, int MODE>
__global__ void __launch_bounds__(GEN_THREADS)
wt_generic_analysis_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ AnaParams P) {
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile_c = blockIdx.x % P.tiles_c;
    const int tile_r = (blockIdx.x / P.tiles_c) % P.tiles_r;
    const long long pg = blockIdx.x / (P.tiles_c * P.tiles_r);
    const long long p0 = pg * P.pb;
    const int np = (int)min((long long)P.pb, P.batch - p0);
    const int o0 = tile_r * P.t2r, c0 = tile_c * P.t2c;
    constexpr int base = (MODE == WT_MODE_PERIODIZATION) ? F / 2 : 1;
    const int r_start = base + 2 * o0 - (F - 1), c_start = base + 2 * c0 - (F - 1);
    const int ir = P.ir, ic = P.ic, ics = P.ics, t2r = P.t2r, t2c = P.t2c;
    const int half = (ic + 1) >> 1, lw = 2 * half;
    const int c_al = (c_start >> 2) << 2;  // floor to a multiple of 4 (arithmetic shift: also for negative starts)
    const int shift = c_start - c_al;      // logical column cc of the tile lives at staged column cc + shift
    float* s_in = smem;                             // [pb][ir][ics]
    float* s_lo = s_in + (size_t)P.pb * ir * ics;    // [pb][t2r][lw]: even columns, then odd columns
    float* s_hi = s_lo + (size_t)P.pb * t2r * lw;
    const int rows_valid = min(t2r, P.h2 - o0), cols_valid = min(t2c, P.w2 - c0);
    // interior tiles of aligned planes are staged with 16-byte copies
    const bool vec = c_al >= 0 && c_al + ics <= P.w && (P.in.rstride & 3) == 0 && (P.in.pstride & 3) == 0 &&
                     (reinterpret_cast<uintptr_t>(P.in.p) & 15u) == 0;

    for (int p = 0; p < np; ++p) {
        const float* __restrict__ plane = P.in.p + (p0 + p) * P.in.pstride;
        for (int rr = warp; rr < ir; rr += NWARP) {
            const float* __restrict__ row = plane + (long long)ext_index<MODE>(r_start + rr, P.h) * P.in.rstride;
            float* dst = s_in + ((size_t)p * ir + rr) * ics;
            if (vec) {
                for (int v = lane; v < (ics >> 2); v += 32) cp_async16(dst + 4 * v, row + c_al + 4 * v);
            } else {
                for (int cc = lane; cc < ics; cc += 32) cp_async4(dst + cc, row + ext_index<MODE>(c_al + cc, P.w));
            }
        }
    }
    cp_async_wait_all();
    __syncthreads();
    // axis -2
    for (int p = 0; p < np; ++p) {
        for (int ol = warp; ol < rows_valid; ol += NWARP) {
            const float* top = s_in + ((size_t)p * ir + 2 * ol + F - 1) * ics + shift;  // tap j reads row 2*ol + F-1-j
            float* lo = s_lo + ((size_t)p * t2r + ol) * lw;
            float* hi = s_hi + ((size_t)p * t2r + ol) * lw;
            for (int cc = lane; cc < ic; cc += 32) {
                float sl = 0.0f, sh = 0.0f;
#pragma unroll
                for (int j = 0; j < F; ++j) {
                    const float v = top[cc - j * ics];
                    sl = WT_MAC(sl, bk.dec_lo[j], v);
                    sh = WT_MAC(sh, bk.dec_hi[j], v);
                }
                const int d = (cc & 1) * half + (cc >> 1);
                lo[d] = sl;
                hi[d] = sh;
            }
        }
    }
    __syncthreads();
    // axis -1: column m = 2*cl + F-1-j lives at [(m & 1) * half + (m >> 1)] = [parity_j * half + cl + shift_j]
    for (int p = 0; p < np; ++p) {
        const long long pl = p0 + p;
        for (int ol = warp; ol < rows_valid; ol += NWARP) {
            const float* rl = s_lo + ((size_t)p * t2r + ol) * lw;
            const float* rh = s_hi + ((size_t)p * t2r + ol) * lw;
            const long long o = o0 + ol;
            for (int cl = lane; cl < cols_valid; cl += 32) {
                float saa = 0.0f, sad = 0.0f, sda = 0.0f, sdd = 0.0f;
#pragma unroll
                for (int j = 0; j < F; ++j) {
                    const int idx = ((F - 1 - j) & 1) * half + cl + ((F - 1 - j) >> 1);
                    const float vl = rl[idx], vh = rh[idx];
                    saa = WT_MAC(saa, bk.dec_lo[j], vl);
                    sad = WT_MAC(sad, bk.dec_hi[j], vl);
                    sda = WT_MAC(sda, bk.dec_lo[j], vh);
                    sdd = WT_MAC(sdd, bk.dec_hi[j], vh);
                }
                const int c = c0 + cl;
                P.aa.p[pl * P.aa.pstride + o * P.aa.rstride + c] = saa;
                P.ad.p[pl * P.ad.pstride + o * P.ad.rstride + c] = sad;
                P.da.p[pl * P.da.pstride + o * P.da.rstride + c] = sda;
                P.dd.p[pl * P.dd.pstride + o * P.dd.rstride + c] = sdd;
            }
        }
    }
}

// ---- large-plane fast path: fixed FX_T2R x 64 output tile, one plane per CTA ----------------------------------------
// Same arithmetic and accumulation order as wt_generic_analysis_kernel; what changes is the instruction count (ncu:
// the runtime-geometry kernel issues ~119 thread-instructions per input element against ~32 no-FMA flops):
// every shared-memory stride is a compile-time constant (LDS with immediate offsets instead of an IMAD + LEA per
// tap), the axis -2 pass computes two output rows per thread from one column of F + 2 staged values (they share
// F - 2 of them), the axis -1 pass computes two adjacent output columns per thread from 64-bit shared loads and
// stores float2s.
// tile heights: measured on 4096 x 4096 db4 (round 2, TMA staging): analysis 32 / 16 / 8 rows -> 115 / 103 / 105 us per
// plane over five levels, synthesis 64 / 32 rows -> 123 / 116.5 us.  The smaller tiles load more halo (38 staged rows for 32
// useful ones) but five CTAs fit an SM instead of two, and the other CTAs' arithmetic covers a CTA's wait for its tile.
#ifndef WT_FX_T2R
#define WT_FX_T2R 16
#endif
#ifndef WT_FS_TR_MAX
#define WT_FS_TR_MAX 32
#endif
constexpr int FX_T2R = WT_FX_T2R, FX_T2C = 64;
constexpr int FX_HALF = 80;          // even (8-byte aligned odd half) and = 16 mod 32 (de-interleaving stores conflict-free)
constexpr int FX_LW = 2 * FX_HALF;
template <int F> struct FxGeom {
    static constexpr int IR = 2 * FX_T2R + F - 2, IC = 2 * FX_T2C + F - 2, ICS = (IC + 3 + 3) & ~3;
    static constexpr int SMEM_FLOATS_ = IR * ICS + 2 * FX_T2R * FX_LW;
    static_assert((IC + 1) / 2 <= FX_HALF, "tile wider than the de-interleaved row");
    static constexpr int EXTRA = 16;   // floats past the tile buffers: the mbarrier, then one nonzero counter per warp
    static_assert(SMEM_FLOATS_ + EXTRA <= SMEM_FLOATS, "fixed tile (+ its mbarrier and counters) does not fit");
};

template <int F, int MODE>
__global__ void __launch_bounds__(GEN_THREADS)
wt_generic_analysis_fixed_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ AnaParams P,
                                 const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(128) float smem[];
    using G = FxGeom<F>;
    constexpr int IR = G::IR, IC = G::IC, ICS = G::ICS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile_c = blockIdx.x % P.tiles_c;
    const int tile_r = (blockIdx.x / P.tiles_c) % P.tiles_r;
    const long long pl = blockIdx.x / (P.tiles_c * P.tiles_r);
    const int o0 = tile_r * FX_T2R, c0 = tile_c * FX_T2C;
    constexpr int base = (MODE == WT_MODE_PERIODIZATION) ? F / 2 : 1;
    const int r_start = base + 2 * o0 - (F - 1), c_start = base + 2 * c0 - (F - 1);
    const int c_al = (c_start >> 2) << 2;
    const int shift = c_start - c_al;
    float* s_in = smem;                 // [IR][ICS]
    float* s_lo = s_in + IR * ICS;      // [T2R][LW]: even columns at [0, HALF), odd columns at [HALF, LW)
    float* s_hi = s_lo + FX_T2R * FX_LW;
    const int rows_valid = min(FX_T2R, P.h2 - o0), cols_valid = min(FX_T2C, P.w2 - c0);
    const bool vec = c_al >= 0 && c_al + ICS <= P.w && (P.in.rstride & 3) == 0 && (P.in.pstride & 3) == 0 &&
                     (reinterpret_cast<uintptr_t>(P.in.p) & 15u) == 0;
    if (P.use_tma) {
        unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + G::SMEM_FLOATS_);   // past the tile buffers
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            mbar_expect(bar, IR * ICS * (unsigned)sizeof(float));
            tma_load_box(s_in, &tmap, c_al, r_start, (int)pl, bar);
        }
        __syncthreads();   // the barrier is initialised for everybody
        mbar_wait(bar, 0);
        if (r_start < 0 || r_start + IR > P.h || c_al < 0 || c_al + ICS > P.w) {   // border tile: the zero fill becomes the extension
            const float* __restrict__ plane = P.in.p + pl * P.in.pstride;
            for (int i = threadIdx.x; i < IR * ICS; i += GEN_THREADS) {
                const int rr = i / ICS, cc = i - rr * ICS;
                const int r = r_start + rr, c = c_al + cc;
                if ((unsigned)r >= (unsigned)P.h || (unsigned)c >= (unsigned)P.w)
                    s_in[i] = plane[(long long)ext_index<MODE>(r, P.h) * P.in.rstride + ext_index<MODE>(c, P.w)];
            }
        }
    } else {
        const float* __restrict__ plane = P.in.p + pl * P.in.pstride;
        for (int rr = warp; rr < IR; rr += NWARP) {
            const float* __restrict__ row = plane + (long long)ext_index<MODE>(r_start + rr, P.h) * P.in.rstride;
            float* dst = s_in + rr * ICS;
            if (vec) {
                for (int v = lane; v < ICS / 4; v += 32) cp_async16(dst + 4 * v, row + c_al + 4 * v);
            } else {
                for (int cc = lane; cc < ICS; cc += 32) cp_async4(dst + cc, row + ext_index<MODE>(c_al + cc, P.w));
            }
        }
        cp_async_wait_all();
    }
    __syncthreads();
    unsigned int* s_nz = reinterpret_cast<unsigned int*>(smem + G::SMEM_FLOATS_ + 2);   // [NWARP], past the mbarrier
    if (P.nz_counts4) {
        // nonzeros of the input (dwt_pruning_NoEntropy.py:58) while the tile is in shared memory: input rows 2*o0 .. and
        // columns 2*c0 .. of this tile -- every input element belongs to exactly one tile -- instead of a pass of its own
        constexpr int OFF = (F - 1) - base;   // staged row / logical column of the tile's first own input row / column
        unsigned int nz = 0;
        const int j = threadIdx.x % (2 * FX_T2C);
        if (2 * c0 + j < P.w)
            for (int i = threadIdx.x / (2 * FX_T2C); i < 2 * FX_T2R && 2 * o0 + i < P.h; i += GEN_THREADS / (2 * FX_T2C))
                nz += (s_in[(i + OFF) * ICS + j + OFF + shift] != 0.0f);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) nz += __shfl_down_sync(0xffffffffu, nz, d);
        if (lane == 0) s_nz[warp] = nz;
    }
    // axis -2: output rows ol and ol + 1 of column cc from staged rows 2*ol .. 2*ol + F + 1
    // (row pair, column) items dealt flat over the CTA: IC is not a multiple of 32
    for (int it = threadIdx.x; it < (FX_T2R / 2) * IC; it += GEN_THREADS) {
        const int ol = 2 * (it / IC), cc = it % IC;
        const float* col = s_in + (2 * ol) * ICS + shift;
        {
            float v[F + 2];
#pragma unroll
            for (int k = 0; k < F + 2; ++k) v[k] = col[k * ICS + cc];
            float sl0 = 0.0f, sh0 = 0.0f, sl1 = 0.0f, sh1 = 0.0f;
#pragma unroll
            for (int j = 0; j < F; ++j) {  // tap j of output row ol + q reads staged row 2*(ol + q) + F-1-j
                sl0 = WT_MAC(sl0, bk.dec_lo[j], v[F - 1 - j]);
                sh0 = WT_MAC(sh0, bk.dec_hi[j], v[F - 1 - j]);
                sl1 = WT_MAC(sl1, bk.dec_lo[j], v[F + 1 - j]);
                sh1 = WT_MAC(sh1, bk.dec_hi[j], v[F + 1 - j]);
            }
            const int d = (cc & 1) * FX_HALF + (cc >> 1);
            s_lo[ol * FX_LW + d] = sl0;
            s_hi[ol * FX_LW + d] = sh0;
            s_lo[(ol + 1) * FX_LW + d] = sl1;
            s_hi[(ol + 1) * FX_LW + d] = sh1;
        }
    }
    __syncthreads();
    if (P.nz_counts4 && threadIdx.x == 0) {
        unsigned int nz = 0;
#pragma unroll
        for (int q = 0; q < NWARP; ++q) nz += s_nz[q];
        if (nz) atomicAdd(reinterpret_cast<unsigned long long*>(P.nz_counts4 + 3), (unsigned long long)nz);
        if (blockIdx.x == 0) P.nz_counts4[0] = P.batch * P.h * P.w;
    }
    // axis -1: output columns cl = 2*lane and cl + 1.  Tile column m = 2*c + F-1-j of output column c lives at
    // [(m & 1) * HALF + (m >> 1)] = [parity_j * HALF + c + ((F-1-j) >> 1)]: both outputs read E[0 .. F/2], O[0 .. F/2]
    constexpr int NV2 = (F / 2 + 2) / 2;  // float2 loads per half row
    const bool st2 = ((reinterpret_cast<uintptr_t>(P.aa.p) | reinterpret_cast<uintptr_t>(P.da.p) |
                       reinterpret_cast<uintptr_t>(P.ad.p) | reinterpret_cast<uintptr_t>(P.dd.p)) & 7u) == 0 &&
                     ((P.aa.rstride | P.da.rstride | P.ad.rstride | P.dd.rstride | P.aa.pstride | P.da.pstride | P.ad.pstride |
                       P.dd.pstride) & 1) == 0;
    const int cl = 2 * lane;
    for (int ol = warp; ol < rows_valid; ol += NWARP) {
        const float2* rl = reinterpret_cast<const float2*>(s_lo + ol * FX_LW) + lane;
        const float2* rh = reinterpret_cast<const float2*>(s_hi + ol * FX_LW) + lane;
        float le[2 * NV2], lo_[2 * NV2], he[2 * NV2], ho[2 * NV2];
#pragma unroll
        for (int i = 0; i < NV2; ++i) {
            const float2 a = rl[i], b = rl[FX_HALF / 2 + i], c = rh[i], d = rh[FX_HALF / 2 + i];
            le[2 * i] = a.x; le[2 * i + 1] = a.y;
            lo_[2 * i] = b.x; lo_[2 * i + 1] = b.y;
            he[2 * i] = c.x; he[2 * i + 1] = c.y;
            ho[2 * i] = d.x; ho[2 * i + 1] = d.y;
        }
        float saa[2] = {0.0f, 0.0f}, sad[2] = {0.0f, 0.0f}, sda[2] = {0.0f, 0.0f}, sdd[2] = {0.0f, 0.0f};
#pragma unroll
        for (int j = 0; j < F; ++j) {
            constexpr int dummy = 0;
            (void)dummy;
            const int par = (F - 1 - j) & 1, sft = (F - 1 - j) >> 1;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const float vl = par ? lo_[q + sft] : le[q + sft];
                const float vh = par ? ho[q + sft] : he[q + sft];
                saa[q] = WT_MAC(saa[q], bk.dec_lo[j], vl);
                sad[q] = WT_MAC(sad[q], bk.dec_hi[j], vl);
                sda[q] = WT_MAC(sda[q], bk.dec_lo[j], vh);
                sdd[q] = WT_MAC(sdd[q], bk.dec_hi[j], vh);
            }
        }
        const long long o = o0 + ol;
        const int c = c0 + cl;
        float* paa = P.aa.p + pl * P.aa.pstride + o * P.aa.rstride + c;
        float* pad = P.ad.p + pl * P.ad.pstride + o * P.ad.rstride + c;
        float* pda = P.da.p + pl * P.da.pstride + o * P.da.rstride + c;
        float* pdd = P.dd.p + pl * P.dd.pstride + o * P.dd.rstride + c;
        if (cl + 1 < cols_valid && st2) {
            *reinterpret_cast<float2*>(paa) = make_float2(saa[0], saa[1]);
            *reinterpret_cast<float2*>(pad) = make_float2(sad[0], sad[1]);
            *reinterpret_cast<float2*>(pda) = make_float2(sda[0], sda[1]);
            *reinterpret_cast<float2*>(pdd) = make_float2(sdd[0], sdd[1]);
        } else {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (cl + q < cols_valid) { paa[q] = saa[q]; pad[q] = sad[q]; pda[q] = sda[q]; pdd[q] = sdd[q]; }
            }
        }
    }
}

struct SynParams {
    WtView aa, da, ad, dd, out;
    long long batch;
    int h2, w2, oh, ow, mode;
    int tr, tc, tiles_r, tiles_c, pb, cr, cc;
    int ccs;  // staged row width of the sub-band tiles (cc + alignment slack, multiple of 4)
    const float* thr_dev;
    float thr_host;
    int use_thr, thr_a;
    long long* counts4;
    int use_tma;  // fixed kernel: the four sub-band tiles come in through the tensor maps
};

// virtual coefficient index and tap parity that produce output sample o
template <int F, int MODE>
__device__ __forceinline__ void syn_index(int o, int* iv, int* par) {
    if (MODE == WT_MODE_PERIODIZATION) {
        constexpr int shift = ((F / 2) % 2 == 0) ? 1 : 0;
        const int q = o - shift;  // >= -1
        *iv = F / 4 + (q >> 1);
        *par = q & 1;
    } else {
        *iv = F / 2 - 1 + (o >> 1);
        *par = o & 1;
    }
}

template <int MODE>
__device__ __forceinline__ int syn_map(int v, int n) {
    if ((unsigned)v < (unsigned)n) return v;
    if (MODE == WT_MODE_PERIODIZATION) return pmodi(v, n);
    return v < 0 ? 0 : n - 1;
}

// a[top - j*stride] is the coefficient paired with tap 2j+par
template <int F, int MODE>
__device__ __forceinline__ float syn_sample(const float* a, const float* d, int stride, int top, int par, const WtBank& bk) {
    constexpr int hf = F / 2;
    if (MODE == WT_MODE_PERIODIZATION) {  // direct accumulation, lo taps then hi taps
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < hf; ++j)
            acc = WT_MAC(acc, par ? bk.rec_lo[2 * j + 1] : bk.rec_lo[2 * j], a[top - j * stride]);
#pragma unroll
        for (int j = 0; j < hf; ++j)
            acc = WT_MAC(acc, par ? bk.rec_hi[2 * j + 1] : bk.rec_hi[2 * j], d[top - j * stride]);
        return acc;
    }
    float sl = 0.0f, sh = 0.0f;
#pragma unroll
    for (int j = 0; j < hf; ++j) sl = WT_MAC(sl, par ? bk.rec_lo[2 * j + 1] : bk.rec_lo[2 * j], a[top - j * stride]);
#pragma unroll
    for (int j = 0; j < hf; ++j) sh = WT_MAC(sh, par ? bk.rec_hi[2 * j + 1] : bk.rec_hi[2 * j], d[top - j * stride]);
    return __fadd_rn(__fadd_rn(0.0f, sl), sh);
}

// One CTA: a tr x tc tile of the reconstructed plane for `pb` planes (idwt2: axis -1 first, then axis -2), the
// threshold applied while the four sub-band tiles are staged, crop and zero count fused into the store.
template <int F, int MODE>
__global__ void __launch_bounds__(GEN_THREADS)
wt_generic_synthesis_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ SynParams P) {
    extern __shared__ float smem[];
    constexpr int hf = F / 2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile_c = blockIdx.x % P.tiles_c;
    const int tile_r = (blockIdx.x / P.tiles_c) % P.tiles_r;
    const long long pg = blockIdx.x / (P.tiles_c * P.tiles_r);
    const long long p0 = pg * P.pb;
    const int np = (int)min((long long)P.pb, P.batch - p0);
    const int R0 = tile_r * P.tr, C0 = tile_c * P.tc;
    const int cr = P.cr, cc = P.ccs, tc = P.tc;  // cc: staged row width
    int iv0, par0;
    syn_index<F, MODE>(R0, &iv0, &par0);
    const int row_lo = iv0 - (hf - 1);
    syn_index<F, MODE>(C0, &iv0, &par0);
    const int col_lo0 = iv0 - (hf - 1);
    const int col_lo = (col_lo0 >> 2) << 2;  // staged region starts 16-byte aligned (floor, also for negative starts)
    const size_t band = (size_t)P.pb * cr * cc;
    float* s_aa = smem;
    float* s_da = s_aa + band;
    float* s_ad = s_da + band;
    float* s_dd = s_ad + band;
    float* s_lo = s_dd + band;  // [pb][cr][tc]
    float* s_hi = s_lo + (size_t)P.pb * cr * tc;
    const float t = P.use_thr ? (P.thr_dev ? __ldg(P.thr_dev) : P.thr_host) : 0.0f;
    const bool thr = P.use_thr != 0, thr_a = thr && P.thr_a != 0;
    const int rows_valid = min(P.tr, P.oh - R0), cols_valid = min(tc, P.ow - C0);

    auto aligned_view = [](const WtView& v) {
        return (reinterpret_cast<uintptr_t>(v.p) & 15u) == 0 && (v.rstride & 3) == 0 && (v.pstride & 3) == 0;
    };
    const bool vec = col_lo >= 0 && col_lo + cc <= P.w2 && aligned_view(P.aa) && aligned_view(P.da) && aligned_view(P.ad) &&
                     aligned_view(P.dd);
    for (int p = 0; p < np; ++p) {
        const long long pl = p0 + p;
        for (int r2 = warp; r2 < cr; r2 += NWARP) {
            const long long sr = syn_map<MODE>(row_lo + r2, P.h2);
            const float* __restrict__ raa = P.aa.p + pl * P.aa.pstride + sr * P.aa.rstride;
            const float* __restrict__ rda = P.da.p + pl * P.da.pstride + sr * P.da.rstride;
            const float* __restrict__ rad = P.ad.p + pl * P.ad.pstride + sr * P.ad.rstride;
            const float* __restrict__ rdd = P.dd.p + pl * P.dd.pstride + sr * P.dd.rstride;
            const size_t off = ((size_t)p * cr + r2) * cc;
            if (vec) {
                for (int v = lane; v < (cc >> 2); v += 32) {
                    cp_async16(s_aa + off + 4 * v, raa + col_lo + 4 * v);
                    cp_async16(s_da + off + 4 * v, rda + col_lo + 4 * v);
                    cp_async16(s_ad + off + 4 * v, rad + col_lo + 4 * v);
                    cp_async16(s_dd + off + 4 * v, rdd + col_lo + 4 * v);
                }
            } else
            for (int c2 = lane; c2 < cc; c2 += 32) {
                const int sc = syn_map<MODE>(col_lo + c2, P.w2);
                cp_async4(s_aa + off + c2, raa + sc);
                cp_async4(s_da + off + c2, rda + sc);
                cp_async4(s_ad + off + c2, rad + sc);
                cp_async4(s_dd + off + c2, rdd + sc);
            }
        }
    }
    cp_async_wait_all();
    if (thr) {  // np.where(|c| < t, 0, c): every thread sweeps exactly the elements it copied (TMA: visible to everybody)
        for (int p = 0; p < np; ++p) {
            for (int r2 = warp; r2 < cr; r2 += NWARP) {
                const size_t off = ((size_t)p * cr + r2) * cc;
                for (int i = lane; i < (vec ? (cc >> 2) : cc); i += 32)
                for (int c2 = vec ? 4 * i : i; c2 < (vec ? 4 * i + 4 : i + 1); ++c2) {
                    if (thr_a && fabsf(s_aa[off + c2]) < t) s_aa[off + c2] = 0.0f;
                    if (fabsf(s_da[off + c2]) < t) s_da[off + c2] = 0.0f;
                    if (fabsf(s_ad[off + c2]) < t) s_ad[off + c2] = 0.0f;
                    if (fabsf(s_dd[off + c2]) < t) s_dd[off + c2] = 0.0f;
                }
            }
        }
    }
    __syncthreads();
    // axis -1 first (idwt2)
    for (int p = 0; p < np; ++p) {
        for (int r2 = warp; r2 < cr; r2 += NWARP) {
            const size_t rowoff = ((size_t)p * cr + r2) * cc;
            const size_t dst = ((size_t)p * cr + r2) * tc;
            for (int cl = lane; cl < cols_valid; cl += 32) {
                int iv, par;
                syn_index<F, MODE>(C0 + cl, &iv, &par);
                const int top = iv - col_lo;
                s_lo[dst + cl] = syn_sample<F, MODE>(s_aa + rowoff, s_ad + rowoff, 1, top, par, bk);
                s_hi[dst + cl] = syn_sample<F, MODE>(s_da + rowoff, s_dd + rowoff, 1, top, par, bk);
            }
        }
    }
    __syncthreads();
    unsigned int zc = 0, wr = 0;
    for (int p = 0; p < np; ++p) {
        float* __restrict__ plane = P.out.p + (p0 + p) * P.out.pstride;
        for (int rl = warp; rl < rows_valid; rl += NWARP) {
            int iv, par;
            syn_index<F, MODE>(R0 + rl, &iv, &par);
            const int top = (iv - row_lo) * tc;
            float* __restrict__ orow = plane + (long long)(R0 + rl) * P.out.rstride + C0;
            const size_t off = (size_t)p * cr * tc;
            for (int cl = lane; cl < cols_valid; cl += 32) {
                const float v = syn_sample<F, MODE>(s_lo + off + cl, s_hi + off + cl, tc, top, par, bk);
                orow[cl] = v;
                zc += (v == 0.0f);
                ++wr;
            }
        }
    }
    if (P.counts4) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            zc += __shfl_down_sync(0xffffffffu, zc, d);
            wr += __shfl_down_sync(0xffffffffu, wr, d);
        }
        if (lane == 0 && wr) {
            atomicAdd(reinterpret_cast<unsigned long long*>(P.counts4 + 2), (unsigned long long)zc);
            atomicAdd(reinterpret_cast<unsigned long long*>(P.counts4 + 1), (unsigned long long)(wr - zc));
        }
    }
}

// ---- large-plane fast path of the synthesis: fixed TR x 128 output tile, one plane per CTA ----------------------
// Same arithmetic and accumulation order as wt_generic_synthesis_kernel.  Output samples 2u and 2u + 1 (in the
// shifted index q of syn_index) read the SAME hf coefficients of each band and differ only in the tap parity, so a
// thread computes both from one set of shared-memory loads with compile-time tap indices (the runtime-geometry
// kernel loads twice and selects every tap by parity), on both axes; the axis -1 result is stored with the two
// parities de-interleaved so that neither pass has a shared-memory bank conflict.
constexpr int FS_TC = 128;
constexpr int FS_HALF = 80;  // >= TC/2 + 1 pairs, = 16 mod 32
constexpr int FS_TW = 2 * FS_HALF;
template <int F, int TR> struct FsGeom {
    static constexpr int hf = F / 2;
    static constexpr int CR = TR / 2 + hf + 2, CC = FS_TC / 2 + hf + 2, CCS = (CC + 3 + 3) & ~3;
    static constexpr int BAND = (CR * CCS + 31) & ~31;   // every sub-band tile starts 128-byte aligned (TMA destination)
    static constexpr int SMEM_FLOATS_ = 4 * BAND + 2 * CR * FS_TW;
};
// tile height of the fixed kernel for filter length F: 64 rows when the tile fits shared memory, else 32
template <int F> constexpr int fs_tr() { return (WT_FS_TR_MAX >= 64 && FsGeom<F, 64>::SMEM_FLOATS_ + 4 <= SMEM_FLOATS) ? 64 : 32; }  // + the mbarrier

// the two outputs (tap parity 0, 1) fed by coefficients a[0], a[-stride], ..., a[-(hf-1)*stride] and the same of d
template <int F, int MODE, int STRIDE>
__device__ __forceinline__ void syn_pair(const float* a, const float* d, const WtBank& bk, float& o0, float& o1) {
    constexpr int hf = F / 2;
    float va[hf], vd[hf];
#pragma unroll
    for (int j = 0; j < hf; ++j) { va[j] = a[-j * STRIDE]; vd[j] = d[-j * STRIDE]; }
    if (MODE == WT_MODE_PERIODIZATION) {
        float e = 0.0f, o = 0.0f;
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            e = WT_MAC(e, bk.rec_lo[2 * j], va[j]);
            o = WT_MAC(o, bk.rec_lo[2 * j + 1], va[j]);
        }
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            e = WT_MAC(e, bk.rec_hi[2 * j], vd[j]);
            o = WT_MAC(o, bk.rec_hi[2 * j + 1], vd[j]);
        }
        o0 = e; o1 = o;
    } else {
        float le = 0.0f, lo = 0.0f, he = 0.0f, ho = 0.0f;
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            le = WT_MAC(le, bk.rec_lo[2 * j], va[j]);
            lo = WT_MAC(lo, bk.rec_lo[2 * j + 1], va[j]);
        }
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            he = WT_MAC(he, bk.rec_hi[2 * j], vd[j]);
            ho = WT_MAC(ho, bk.rec_hi[2 * j + 1], vd[j]);
        }
        o0 = __fadd_rn(__fadd_rn(0.0f, le), he);
        o1 = __fadd_rn(__fadd_rn(0.0f, lo), ho);
    }
}

// syn_pair on values already in registers: va[j] / vd[j] = the coefficients paired with taps 2j, 2j + 1 (same order of operations)
template <int F, int MODE>
__device__ __forceinline__ void syn_pair_regs(const float (&va)[F / 2], const float (&vd)[F / 2], const WtBank& bk, float& o0, float& o1) {
    constexpr int hf = F / 2;
    if (MODE == WT_MODE_PERIODIZATION) {
        float e = 0.0f, o = 0.0f;
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            e = WT_MAC(e, bk.rec_lo[2 * j], va[j]);
            o = WT_MAC(o, bk.rec_lo[2 * j + 1], va[j]);
        }
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            e = WT_MAC(e, bk.rec_hi[2 * j], vd[j]);
            o = WT_MAC(o, bk.rec_hi[2 * j + 1], vd[j]);
        }
        o0 = e; o1 = o;
    } else {
        float le = 0.0f, lo = 0.0f, he = 0.0f, ho = 0.0f;
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            le = WT_MAC(le, bk.rec_lo[2 * j], va[j]);
            lo = WT_MAC(lo, bk.rec_lo[2 * j + 1], va[j]);
        }
#pragma unroll
        for (int j = 0; j < hf; ++j) {
            he = WT_MAC(he, bk.rec_hi[2 * j], vd[j]);
            ho = WT_MAC(ho, bk.rec_hi[2 * j + 1], vd[j]);
        }
        o0 = __fadd_rn(__fadd_rn(0.0f, le), he);
        o1 = __fadd_rn(__fadd_rn(0.0f, lo), ho);
    }
}

template <int F, int MODE, int TR>
__global__ void __launch_bounds__(GEN_THREADS)
wt_generic_synthesis_fixed_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ SynParams P,
                                  const __grid_constant__ CUtensorMap map_aa, const __grid_constant__ CUtensorMap map_da,
                                  const __grid_constant__ CUtensorMap map_ad, const __grid_constant__ CUtensorMap map_dd) {
    extern __shared__ __align__(128) float smem[];
    using G = FsGeom<F, TR>;
    constexpr int hf = F / 2, CR = G::CR, CCS = G::CCS;
    // q = o - SH is the index whose parity picks the taps (syn_index)
    constexpr int SH = (MODE == WT_MODE_PERIODIZATION && (hf % 2 == 0)) ? 1 : 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile_c = blockIdx.x % P.tiles_c;
    const int tile_r = (blockIdx.x / P.tiles_c) % P.tiles_r;
    const long long pl = blockIdx.x / (P.tiles_c * P.tiles_r);
    const int R0 = tile_r * TR, C0 = tile_c * FS_TC;
    int iv0, par0;
    syn_index<F, MODE>(R0, &iv0, &par0);
    const int row_lo = iv0 - (hf - 1);
    syn_index<F, MODE>(C0, &iv0, &par0);
    const int col_lo0 = iv0 - (hf - 1);
    const int col_lo = (col_lo0 >> 2) << 2;
    const int d0 = col_lo0 - col_lo;
    constexpr int band = G::BAND;
    float* s_aa = smem;
    float* s_da = s_aa + band;
    float* s_ad = s_da + band;
    float* s_dd = s_ad + band;
    float* s_lo = s_dd + band;  // [CR][TW]: pair t of parity p at [p * HALF + t]
    float* s_hi = s_lo + CR * FS_TW;
    const float t = P.use_thr ? (P.thr_dev ? __ldg(P.thr_dev) : P.thr_host) : 0.0f;
    const bool thr = P.use_thr != 0, thr_a = thr && P.thr_a != 0;
    const int rows_valid = min(TR, P.oh - R0), cols_valid = min(FS_TC, P.ow - C0);

    auto aligned_view = [](const WtView& v) {
        return (reinterpret_cast<uintptr_t>(v.p) & 15u) == 0 && (v.rstride & 3) == 0 && (v.pstride & 3) == 0;
    };
    const bool vec = col_lo >= 0 && col_lo + CCS <= P.w2 && aligned_view(P.aa) && aligned_view(P.da) && aligned_view(P.ad) &&
                     aligned_view(P.dd);
    const bool tma = P.use_tma != 0;
    if (tma) {
        unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + G::SMEM_FLOATS_);   // past the tile buffers
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            mbar_expect(bar, 4u * CR * CCS * (unsigned)sizeof(float));
            tma_load_box(s_aa, &map_aa, col_lo, row_lo, (int)pl, bar);
            tma_load_box(s_da, &map_da, col_lo, row_lo, (int)pl, bar);
            tma_load_box(s_ad, &map_ad, col_lo, row_lo, (int)pl, bar);
            tma_load_box(s_dd, &map_dd, col_lo, row_lo, (int)pl, bar);
        }
        __syncthreads();   // the barrier is initialised for everybody
        mbar_wait(bar, 0);
        if (row_lo < 0 || row_lo + CR > P.h2 || col_lo < 0 || col_lo + CCS > P.w2) {   // border tile: zero fill -> extension
            for (int i = threadIdx.x; i < CR * CCS; i += GEN_THREADS) {
                const int r2 = i / CCS, c2 = i - r2 * CCS;
                const int r = row_lo + r2, c = col_lo + c2;
                if ((unsigned)r >= (unsigned)P.h2 || (unsigned)c >= (unsigned)P.w2) {
                    const long long sr = syn_map<MODE>(r, P.h2);
                    const int sc = syn_map<MODE>(c, P.w2);
                    s_aa[i] = P.aa.p[pl * P.aa.pstride + sr * P.aa.rstride + sc];
                    s_da[i] = P.da.p[pl * P.da.pstride + sr * P.da.rstride + sc];
                    s_ad[i] = P.ad.p[pl * P.ad.pstride + sr * P.ad.rstride + sc];
                    s_dd[i] = P.dd.p[pl * P.dd.pstride + sr * P.dd.rstride + sc];
                }
            }
            __syncthreads();   // the threshold sweep below reads what other threads patched
        }
    } else {
        for (int r2 = warp; r2 < CR; r2 += NWARP) {
            const long long sr = syn_map<MODE>(row_lo + r2, P.h2);
            const float* __restrict__ raa = P.aa.p + pl * P.aa.pstride + sr * P.aa.rstride;
            const float* __restrict__ rda = P.da.p + pl * P.da.pstride + sr * P.da.rstride;
            const float* __restrict__ rad = P.ad.p + pl * P.ad.pstride + sr * P.ad.rstride;
            const float* __restrict__ rdd = P.dd.p + pl * P.dd.pstride + sr * P.dd.rstride;
            const int off = r2 * CCS;
            if (vec) {
                for (int v = lane; v < CCS / 4; v += 32) {
                    cp_async16(s_aa + off + 4 * v, raa + col_lo + 4 * v);
                    cp_async16(s_da + off + 4 * v, rda + col_lo + 4 * v);
                    cp_async16(s_ad + off + 4 * v, rad + col_lo + 4 * v);
                    cp_async16(s_dd + off + 4 * v, rdd + col_lo + 4 * v);
                }
            } else {
                for (int c2 = lane; c2 < CCS; c2 += 32) {
                    const int sc = syn_map<MODE>(col_lo + c2, P.w2);
                    cp_async4(s_aa + off + c2, raa + sc);
                    cp_async4(s_da + off + c2, rda + sc);
                    cp_async4(s_ad + off + c2, rad + sc);
                    cp_async4(s_dd + off + c2, rdd + sc);
                }
            }
        }
    
        cp_async_wait_all();
    }
    if (thr) {  // np.where(|c| < t, 0, c): every thread sweeps exactly the elements it copied
        auto cut = [t](float v) { return fabsf(v) < t ? 0.0f : v; };
        auto cut4 = [&](float* p) {   // 16-byte shared accesses: a lane's four floats in one conflict-free wavefront
            float4 v = *reinterpret_cast<float4*>(p);
            v.x = cut(v.x); v.y = cut(v.y); v.z = cut(v.z); v.w = cut(v.w);
            *reinterpret_cast<float4*>(p) = v;
        };
        for (int r2 = warp; r2 < CR; r2 += NWARP) {
            const int off = r2 * CCS;
            if (vec || tma) {
                for (int i = lane; i < CCS / 4; i += 32) {
                    if (thr_a) cut4(s_aa + off + 4 * i);
                    cut4(s_da + off + 4 * i);
                    cut4(s_ad + off + 4 * i);
                    cut4(s_dd + off + 4 * i);
                }
            } else {
                for (int c2 = lane; c2 < CCS; c2 += 32) {
                    if (thr_a) s_aa[off + c2] = cut(s_aa[off + c2]);
                    s_da[off + c2] = cut(s_da[off + c2]);
                    s_ad[off + c2] = cut(s_ad[off + c2]);
                    s_dd[off + c2] = cut(s_dd[off + c2]);
                }
            }
        }
    }
    __syncthreads();
    // axis -1 first (idwt2).  Pair tp: q = 2*(tp - SH), q + 1, i.e. tile columns 2*tp - SH and 2*tp - SH + 1; both read
    // the staged coefficients [tp + hf-1 + d0 - j], j < hf
    constexpr int NPAIR_C = FS_TC / 2 + SH;
    constexpr int ROWS_USED = (TR / 2 + SH) + hf - 1;   // intermediate rows the axis -2 pass reads (CR keeps two spare)
    static_assert(ROWS_USED <= CR, "staged rows");
    // (row, pair) items dealt flat over the CTA: neither ROWS_USED nor NPAIR_C is a multiple of the warp geometry
    for (int it = threadIdx.x; it < ROWS_USED * NPAIR_C; it += GEN_THREADS) {
        const int r2 = it / NPAIR_C, tp = it % NPAIR_C;
        const float* aa = s_aa + r2 * CCS + (hf - 1) + d0;
        const float* ad = s_ad + r2 * CCS + (hf - 1) + d0;
        const float* da = s_da + r2 * CCS + (hf - 1) + d0;
        const float* dd = s_dd + r2 * CCS + (hf - 1) + d0;
        {
            float l0, l1, h0, h1;
            syn_pair<F, MODE, 1>(aa + tp, ad + tp, bk, l0, l1);
            syn_pair<F, MODE, 1>(da + tp, dd + tp, bk, h0, h1);
            s_lo[r2 * FS_TW + tp] = l0;
            s_lo[r2 * FS_TW + FS_HALF + tp] = l1;
            s_hi[r2 * FS_TW + tp] = h0;
            s_hi[r2 * FS_TW + FS_HALF + tp] = h1;
        }
    }
    __syncthreads();
    // axis -2: row pair tp (tile rows 2*tp - SH, 2*tp - SH + 1) of column cl reads intermediate rows tp + hf-1 - j.
    // A thread walks DOWN its column: consecutive row pairs share hf - 1 of their hf intermediate rows, so the window lives
    // in registers and a step costs two shared loads (one new row of s_lo, one of s_hi) and two stores at a running pointer
    // (ncu, round 2: the flat (pair, column) loop spent 86 instructions per pair on 31 flops -- eight loads, the address
    // arithmetic of both stores and the row checks, every pair from scratch).
    unsigned int zc = 0, wr = 0;
    constexpr int NPAIR_R = TR / 2 + SH;
    constexpr int COL_GROUPS = GEN_THREADS / FS_TC;                 // threads per column
    constexpr int PAIRS_PER = (NPAIR_R + COL_GROUPS - 1) / COL_GROUPS;
    static_assert(GEN_THREADS % (2 * FX_T2C) == 0, "whole input columns per thread group");
    static_assert(GEN_THREADS % FS_TC == 0, "whole columns per thread group");
    {
        const int cl = threadIdx.x % FS_TC, grp = threadIdx.x / FS_TC;
        const int tp0 = grp * PAIRS_PER, tp1 = min(NPAIR_R, tp0 + PAIRS_PER);
        if (cl < cols_valid && tp0 < tp1) {
            const int idx = ((cl + SH) & 1) * FS_HALF + ((cl + SH) >> 1);
            const float* lo = s_lo + idx;
            const float* hi = s_hi + idx;
            float wl[hf], wh[hf];   // w[j] = intermediate row tp + hf-1 - j
#pragma unroll
            for (int j = 1; j < hf; ++j) {
                wl[j - 1] = lo[(tp0 + hf - 1 - j) * FS_TW];
                wh[j - 1] = hi[(tp0 + hf - 1 - j) * FS_TW];
            }
            float* __restrict__ o = P.out.p + pl * P.out.pstride + C0 + (long long)(R0 + 2 * tp0 - SH) * P.out.rstride + cl;
            const long long rs = P.out.rstride;
            for (int tp = tp0; tp < tp1; ++tp) {
#pragma unroll
                for (int j = hf - 1; j > 0; --j) { wl[j] = wl[j - 1]; wh[j] = wh[j - 1]; }
                wl[0] = lo[(tp + hf - 1) * FS_TW];
                wh[0] = hi[(tp + hf - 1) * FS_TW];
                float v0, v1;
                syn_pair_regs<F, MODE>(wl, wh, bk, v0, v1);
                const int rl0 = 2 * tp - SH;
                if (rl0 >= 0 && rl0 < rows_valid) { o[0] = v0; zc += (v0 == 0.0f); ++wr; }
                if (rl0 + 1 < rows_valid) { o[rs] = v1; zc += (v1 == 0.0f); ++wr; }
                o += 2 * rs;
            }
        }
    }
    if (P.counts4) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            zc += __shfl_down_sync(0xffffffffu, zc, d);
            wr += __shfl_down_sync(0xffffffffu, wr, d);
        }
        if (lane == 0 && wr) {
            atomicAdd(reinterpret_cast<unsigned long long*>(P.counts4 + 2), (unsigned long long)zc);
            atomicAdd(reinterpret_cast<unsigned long long*>(P.counts4 + 1), (unsigned long long)(wr - zc));
        }
    }
}

__global__ void __launch_bounds__(256) wt_count_nonzero_kernel(const float* __restrict__ x, long long n,
                                                               long long* counts4) {
    unsigned int nz = 0;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
    long long done = 0;
    if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
        const long long nvec = n >> 2;
        const float4* __restrict__ x4 = reinterpret_cast<const float4*>(x);
        for (long long v = tid; v < nvec; v += 4 * nthr) {
            float4 r[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (v + u * nthr < nvec) r[u] = __ldg(&x4[v + u * nthr]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (v + u * nthr < nvec) nz += (r[u].x != 0.0f) + (r[u].y != 0.0f) + (r[u].z != 0.0f) + (r[u].w != 0.0f);
        }
        done = nvec << 2;
    }
    for (long long i = done + tid; i < n; i += nthr) nz += (x[i] != 0.0f);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) nz += __shfl_down_sync(0xffffffffu, nz, d);
    if ((threadIdx.x & 31) == 0 && nz) atomicAdd(reinterpret_cast<unsigned long long*>(counts4 + 3), (unsigned long long)nz);
    if (blockIdx.x == 0 && threadIdx.x == 0) counts4[0] = n;
}

// ---- the other seven pywt modes (zero, constant, reflect, periodic, smooth, antisymmetric, antireflect) -------------
// pywt.dwt2 is two dwt_axis sweeps (axis -2, then axis -1 over both results) and the extension of the second sweep is
// formed from the FIRST sweep's output, which for smooth / antireflect is not what a 2-D extended input tile would give
// in fp32.  These modes are off the reference's default path (dwt_pruning.py:35 defaults to 'periodization'), so they
// take a plain two-sweep form through a scratch buffer: one thread per output sample, the extended sample is formed
// first (index map, negation, or subtract-multiply-add), then multiplied by the tap -- operation for operation what
// oracle/dwt_oracle.c does.  Synthesis needs nothing new: pywt's idwt ignores the mode unless it is periodization.
__device__ __forceinline__ float antireflect_side(const float* __restrict__ x, int n, long long xs, int d, int near_i, int far_i) {
    const float xn = x[near_i * xs], xf = x[far_i * xs];
    const int step = far_i > near_i ? 1 : -1;
    float le = xn;
    int blk = 0;
    while (d > n - 1) {
        const float t = (blk & 1) ? __fsub_rn(xn, xf) : __fsub_rn(xf, xn);
        le = (blk & 1) ? __fadd_rn(le, t) : __fsub_rn(le, t);
        d -= n - 1;
        ++blk;
    }
    if (blk & 1) return __fadd_rn(le, __fsub_rn(x[(far_i - step * d) * xs], xf));
    return __fsub_rn(le, __fsub_rn(x[(near_i + step * d) * xs], xn));
}

__device__ __forceinline__ float ext_value(const float* __restrict__ x, int n, long long xs, int idx, int mode) {
    if ((unsigned)idx < (unsigned)n) return x[idx * xs];
    switch (mode) {
    case WT_MODE_ZERO:
        return 0.0f;
    case WT_MODE_CONSTANT:
        return x[(idx < 0 ? 0 : n - 1) * xs];
    case WT_MODE_REFLECT: {
        if (n == 1) return x[0];
        const int per = 2 * n - 2;
        int m = pmodi(idx, per);
        if (m >= n) m = per - m;
        return x[m * xs];
    }
    case WT_MODE_PERIODIC:
        return x[pmodi(idx, n) * xs];
    case WT_MODE_SMOOTH: {
        if (n < 2) return x[0];
        if (idx < 0) {
            const float slope = __fsub_rn(x[0], x[xs]);
            return __fadd_rn(x[0], __fmul_rn(slope, (float)(-idx)));
        }
        const float e = x[(n - 1) * xs];
        const float slope = __fsub_rn(e, x[(n - 2) * xs]);
        return __fadd_rn(e, __fmul_rn(slope, (float)(idx - (n - 1))));
    }
    case WT_MODE_ANTISYMMETRIC: {
        const int m = pmodi(idx, 2 * n);
        return m < n ? x[m * xs] : -x[(2 * n - 1 - m) * xs];
    }
    case WT_MODE_ANTIREFLECT:
        if (n == 1) return x[0];
        return idx < 0 ? antireflect_side(x, n, xs, -idx, 0, n - 1) : antireflect_side(x, n, xs, idx - (n - 1), n - 1, 0);
    default: {  // symmetric
        int m = pmodi(idx, 2 * n);
        if (m >= n) m = 2 * n - 1 - m;
        return x[m * xs];
    }
    }
}

struct AxisParams {
    WtView in, aa, da, ad, dd;
    float* lo;  // [batch][h2][w] results of the axis -2 sweep
    float* hi;
    long long batch;
    int h, w, h2, w2, mode, F;
};

// axis -2: in (h x w) -> lo, hi (h2 x w); consecutive threads walk a row, so the F reads of a warp are coalesced
__global__ void __launch_bounds__(256)
wt_generic_axis0_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ AxisParams P) {
    const long long total = P.batch * P.h2 * P.w;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += 256LL * gridDim.x) {
        const int c = (int)(i % P.w);
        const long long t = i / P.w;
        const int o = (int)(t % P.h2);
        const long long p = t / P.h2;
        const float* __restrict__ col = P.in.p + p * P.in.pstride + c;
        float sl = 0.0f, sh = 0.0f;
        for (int j = 0; j < P.F; ++j) {
            const float v = ext_value(col, P.h, P.in.rstride, 1 + 2 * o - j, P.mode);
            sl = WT_MAC(sl, bk.dec_lo[j], v);
            sh = WT_MAC(sh, bk.dec_hi[j], v);
        }
        P.lo[i] = sl;
        P.hi[i] = sh;
    }
}

// axis -1: lo, hi (h2 x w) -> aa, ad (from lo) and da, dd (from hi), each h2 x w2
__global__ void __launch_bounds__(256)
wt_generic_axis1_kernel(const __grid_constant__ WtBank bk, const __grid_constant__ AxisParams P) {
    const long long total = P.batch * P.h2 * P.w2;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += 256LL * gridDim.x) {
        const int c = (int)(i % P.w2);
        const long long t = i / P.w2;
        const int o = (int)(t % P.h2);
        const long long p = t / P.h2;
        const float* __restrict__ rl = P.lo + t * P.w;
        const float* __restrict__ rh = P.hi + t * P.w;
        float saa = 0.0f, sad = 0.0f, sda = 0.0f, sdd = 0.0f;
        for (int j = 0; j < P.F; ++j) {
            const float vl = ext_value(rl, P.w, 1, 1 + 2 * c - j, P.mode);
            const float vh = ext_value(rh, P.w, 1, 1 + 2 * c - j, P.mode);
            saa = WT_MAC(saa, bk.dec_lo[j], vl);
            sad = WT_MAC(sad, bk.dec_hi[j], vl);
            sda = WT_MAC(sda, bk.dec_lo[j], vh);
            sdd = WT_MAC(sdd, bk.dec_hi[j], vh);
        }
        P.aa.p[p * P.aa.pstride + (long long)o * P.aa.rstride + c] = saa;
        P.ad.p[p * P.ad.pstride + (long long)o * P.ad.rstride + c] = sad;
        P.da.p[p * P.da.pstride + (long long)o * P.da.rstride + c] = sda;
        P.dd.p[p * P.dd.pstride + (long long)o * P.dd.rstride + c] = sdd;
    }
}

}  // namespace

// filter length x mode -> template instantiation; WT_GEN_LAUNCH(F, MODE) is defined at the point of use
#define WT_GEN_CASE(FL)                                                  \
    case FL:                                                             \
        if (mode_ == WT_MODE_PERIODIZATION) WT_GEN_LAUNCH(FL, WT_MODE_PERIODIZATION) \
        else WT_GEN_LAUNCH(FL, WT_MODE_SYMMETRIC)                        \
        break;
#define WT_GEN_DISPATCH(F_, mode__)                                                            \
    {                                                                                          \
        const int mode_ = (mode__);                                                            \
        switch (F_) {                                                                          \
            WT_GEN_CASE(2) WT_GEN_CASE(4) WT_GEN_CASE(6) WT_GEN_CASE(8) WT_GEN_CASE(10) WT_GEN_CASE(12) WT_GEN_CASE(14) WT_GEN_CASE(16) WT_GEN_CASE(18) \
            default:                                                                           \
                wt_set_error("generic path: unsupported filter length %d", (int)(F_));         \
                return WT_E_UNSUPPORTED;                                                       \
        }                                                                                      \
    }

// ---- host side -----------------------------------------------------------------------------------------
#ifndef WT_GENERIC_FMA
long long wt_generic_ncoef(int H, int W, int F, int mode, int level) {
    int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1];
    wt_level_sizes(H, F, mode, level, hs);
    wt_level_sizes(W, F, mode, level, ws);
    long long n = (long long)hs[level] * ws[level];
    for (int l = 1; l <= level; ++l) n += 3LL * hs[l] * ws[l];
    return n;
}

// per-plane floats of one ping-pong buffer: [approximation / reconstruction | scratch of the two-sweep modes]
static long long tmp_pingpong_elems(int H, int W, int F, int mode, int level) {
    int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1];
    wt_level_sizes(H, F, mode, level, hs);
    wt_level_sizes(W, F, mode, level, ws);
    long long m = 1;
    for (int l = 1; l <= level; ++l) {
        m = std::max(m, (long long)hs[l] * ws[l]);                                                   // analysis cA_l
        m = std::max(m, (long long)std::max(hs[l - 1], 1) * std::max(ws[l - 1], 1));                  // synthesis output
    }
    return m;
}

long long wt_generic_tmp_elems(int H, int W, int F, int mode, int level) {
    long long m = tmp_pingpong_elems(H, W, F, mode, level);
    if (mode > WT_MODE_SYMMETRIC) {  // lo + hi of the axis -2 sweep, behind the ping-pong part
        int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1];
        wt_level_sizes(H, F, mode, level, hs);
        wt_level_sizes(W, F, mode, level, ws);
        long long sc = 0;
        for (int l = 1; l <= level; ++l) sc = std::max(sc, 2LL * hs[l] * ws[l - 1]);
        m += sc;
    }
    return m;
}

float* wt_generic_scratch(float* tmp, long long batch, int H, int W, int F, int mode, int level) {
    return mode > WT_MODE_SYMMETRIC ? tmp + batch * tmp_pingpong_elems(H, W, F, mode, level) : nullptr;
}

#endif  // !WT_GENERIC_FMA

// ---- tensor maps for the fixed-geometry kernels ----------------------------------------------------------------
// cuTensorMapEncodeTiled lives in the driver library; the entry point comes from the runtime, so libwtprune.so keeps
// linking against cudart alone.
typedef CUresult (*WtEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static WtEncodeTiled encode_tiled_entry() {
    static WtEncodeTiled fn = []() -> WtEncodeTiled {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<WtEncodeTiled>(p);
    }();
    return fn;
}

// map of the view as a 3-D tensor {w, h, batch} with a box of box_rows x box_cols floats of one plane; false when the
// view cannot be described (unaligned base / strides): the kernels then stage with cp.async as before
static bool make_plane_map(CUtensorMap* m, const WtView& v, long long batch, int h, int w, int box_rows, int box_cols) {
    memset(m, 0, sizeof(*m));
    static const bool off = getenv("WT_GENERIC_NO_TMA") != nullptr;
    WtEncodeTiled fn = off ? nullptr : encode_tiled_entry();
    if (!fn) return false;
    if ((reinterpret_cast<uintptr_t>(v.p) & 15u) != 0 || (v.rstride & 3) != 0 || (batch > 1 && (v.pstride & 3) != 0)) return false;
    if (box_cols > 256 || box_rows > 256 || (box_cols & 3) != 0 || h <= 0 || w <= 0 || batch <= 0 || batch > 2147483647LL) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)batch};
    const cuuint64_t pstride_b = (cuuint64_t)(batch > 1 ? v.pstride : (long long)v.rstride * h) * sizeof(float);
    const cuuint64_t strides[2] = {(cuuint64_t)v.rstride * sizeof(float), (pstride_b + 15) / 16 * 16};
    const cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, v.p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int WT_GEN_FN(wt_generic_analysis_level)(const WtBank& bank, int mode, WtView in, long long batch, int h, int w, WtView aa,
                              WtView da, WtView ad, WtView dd, float* scratch, cudaStream_t stream, long long* nz_counts4,
                              int* nz_counted) {
    if (nz_counted) *nz_counted = 0;
    if (mode > WT_MODE_SYMMETRIC) {
        WT_CHECK(scratch != nullptr, WT_E_INVALID, "generic analysis: mode %d needs the sweep scratch", mode);
        AxisParams A;
        A.in = in; A.aa = aa; A.da = da; A.ad = ad; A.dd = dd;
        A.batch = batch; A.h = h; A.w = w; A.mode = mode; A.F = bank.F;
        A.h2 = wt_dwt_coeff_len(h, bank.F, mode);
        A.w2 = wt_dwt_coeff_len(w, bank.F, mode);
        A.lo = scratch;
        A.hi = scratch + batch * A.h2 * w;
        const long long n0 = batch * A.h2 * w, n1 = batch * A.h2 * A.w2;
        if (n0 <= 0 || n1 <= 0) return WT_OK;
        const long long cap = 148LL * 8 * 4;
        wt_generic_axis0_kernel<<<(unsigned)std::min((n0 + 255) / 256, cap), 256, 0, stream>>>(bank, A);
        wt_generic_axis1_kernel<<<(unsigned)std::min((n1 + 255) / 256, cap), 256, 0, stream>>>(bank, A);
        WT_CUDA(cudaGetLastError());
        return WT_OK;
    }
    AnaParams P;
    P.in = in; P.aa = aa; P.da = da; P.ad = ad; P.dd = dd;
    P.batch = batch; P.h = h; P.w = w; P.mode = mode; P.use_tma = 0; P.nz_counts4 = nullptr;
    P.h2 = wt_dwt_coeff_len(h, bank.F, mode);
    P.w2 = wt_dwt_coeff_len(w, bank.F, mode);
    P.t2r = std::min(P.h2, 32);
    P.t2c = std::min(P.w2, 64);
    while ((long long)((P.h2 + P.t2r - 1) / P.t2r) * ((P.w2 + P.t2c - 1) / P.t2c) * batch < 296 && (P.t2r > 4 || P.t2c > 32)) {
        if (P.t2c > 32 && P.t2c >= 2 * P.t2r) P.t2c /= 2; else if (P.t2r > 4) P.t2r /= 2; else P.t2c /= 2;
    }
    P.tiles_r = (P.h2 + P.t2r - 1) / P.t2r;
    P.tiles_c = (P.w2 + P.t2c - 1) / P.t2c;
    P.ir = 2 * P.t2r + bank.F - 2;
    P.ic = 2 * P.t2c + bank.F - 2;
    P.ics = (P.ic + 3 + 3) & ~3;
    const long long per_plane = (long long)P.ir * P.ics + 2LL * P.t2r * (2 * ((P.ic + 1) / 2));
    long long pb = SMEM_FLOATS / per_plane;
    WT_CHECK(pb >= 1, WT_E_UNSUPPORTED, "generic analysis tile does not fit shared memory");
    pb = std::min(pb, batch);
    // keep the CTA busy but do not serialise too many planes behind 256 threads
    pb = std::min(pb, std::max(1LL, 2048LL / ((long long)P.t2r * P.t2c)));
    P.pb = (int)pb;
    const long long groups = (batch + pb - 1) / pb;
    const long long grid = groups * P.tiles_r * P.tiles_c;
    WT_CHECK(grid < 2147483647LL, WT_E_UNSUPPORTED, "generic analysis grid too large");
    const size_t smem = (size_t)pb * per_plane * sizeof(float);
    // large planes: the fixed-geometry kernel (full-size tiles, one plane per CTA)
    const long long ftiles = (long long)((P.h2 + FX_T2R - 1) / FX_T2R) * ((P.w2 + FX_T2C - 1) / FX_T2C);
    const bool fixed = P.h2 >= FX_T2R && P.w2 >= FX_T2C && ftiles * batch >= 296 && !getenv("WT_GENERIC_NO_FIXED");
    if (fixed) {
        if (nz_counts4 && nz_counted && in.rstride == w && in.pstride == (long long)h * w) {   // contiguous input: count it on the way
            P.nz_counts4 = nz_counts4;
            *nz_counted = 1;
        }
        P.t2r = FX_T2R; P.t2c = FX_T2C; P.pb = 1;
        P.tiles_r = (P.h2 + FX_T2R - 1) / FX_T2R;
        P.tiles_c = (P.w2 + FX_T2C - 1) / FX_T2C;
    }
#define WT_GEN_LAUNCH(FL, MD)                                                                                     \
    if (fixed) {                                                                                                  \
        auto kern = wt_generic_analysis_fixed_kernel<FL, MD>;                                                     \
        WT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_FLOATS * 4));        \
        CUtensorMap tm;                                                                                           \
        P.use_tma = make_plane_map(&tm, in, batch, h, w, FxGeom<FL>::IR, FxGeom<FL>::ICS) ? 1 : 0;                \
        kern<<<(unsigned)(ftiles * batch), GEN_THREADS, (FxGeom<FL>::SMEM_FLOATS_ + FxGeom<FL>::EXTRA) * sizeof(float), stream>>>(bank, P, tm); \
    } else {                                                                                                      \
        auto kern = wt_generic_analysis_kernel<FL, MD>;                                                           \
        WT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_FLOATS * 4));        \
        kern<<<(unsigned)grid, GEN_THREADS, smem, stream>>>(bank, P);                                             \
    }
    WT_GEN_DISPATCH(bank.F, mode)
#undef WT_GEN_LAUNCH
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

int WT_GEN_FN(wt_generic_synthesis_level)(const WtBank& bank, int mode, WtView aa, WtView da, WtView ad, WtView dd,
                               long long batch, int h2, int w2, WtView out, int oh, int ow, const float* thr_dev,
                               float thr_host, int use_thr, int thr_a, long long* counts4, cudaStream_t stream) {
    SynParams P;
    P.aa = aa; P.da = da; P.ad = ad; P.dd = dd; P.out = out;
    P.batch = batch; P.h2 = h2; P.w2 = w2; P.oh = oh; P.ow = ow; P.mode = mode;
    P.tr = std::min(oh, 64);
    P.tc = std::min(ow, 128);
    P.tiles_r = (oh + P.tr - 1) / P.tr;
    P.tiles_c = (ow + P.tc - 1) / P.tc;
    // small levels: shrink the tile until the grid can fill the GPU (a 512 x 512 level is only 32 big tiles)
    while ((long long)((oh + P.tr - 1) / P.tr) * ((ow + P.tc - 1) / P.tc) * batch < 296 && (P.tr > 8 || P.tc > 32)) {
        if (P.tc > 32 && P.tc >= 2 * P.tr) P.tc /= 2; else if (P.tr > 8) P.tr /= 2; else P.tc /= 2;
    }
    auto geometry = [&]() {
        P.tiles_r = (oh + P.tr - 1) / P.tr;
        P.tiles_c = (ow + P.tc - 1) / P.tc;
        P.cr = P.tr / 2 + bank.F / 2 + 2;
        P.cc = P.tc / 2 + bank.F / 2 + 2;
        P.ccs = (P.cc + 3 + 3) & ~3;
        return 4LL * P.cr * P.ccs + 2LL * P.cr * P.tc;
    };
    // the longest banks (F = 18) do not fit a full 64 x 128 tile next to its four staged sub-band tiles
    while (geometry() > SMEM_FLOATS && (P.tr > 8 || P.tc > 32)) {
        if (P.tr > 8) P.tr /= 2; else P.tc /= 2;
    }
    P.thr_dev = thr_dev; P.thr_host = thr_host; P.use_thr = use_thr; P.thr_a = thr_a; P.counts4 = counts4; P.use_tma = 0;
    const long long per_plane = 4LL * P.cr * P.ccs + 2LL * P.cr * P.tc;
    long long pb = SMEM_FLOATS / per_plane;
    WT_CHECK(pb >= 1, WT_E_UNSUPPORTED, "generic synthesis tile does not fit shared memory");
    pb = std::min(pb, batch);
    pb = std::min(pb, std::max(1LL, 2048LL / ((long long)P.tr * P.tc)));
    P.pb = (int)pb;
    const long long groups = (batch + pb - 1) / pb;
    const long long grid = groups * P.tiles_r * P.tiles_c;
    WT_CHECK(grid < 2147483647LL, WT_E_UNSUPPORTED, "generic synthesis grid too large");
    const size_t smem = (size_t)pb * per_plane * sizeof(float);
    // large planes: the fixed-geometry kernel (full-size tiles, one plane per CTA)
    const bool no_fixed = getenv("WT_GENERIC_NO_FIXED") != nullptr;
#define WT_GEN_LAUNCH(FL, MD)                                                                                     \
    {                                                                                                             \
        constexpr int FTR = fs_tr<FL>();                                                                          \
        const long long ftiles = (long long)((oh + FTR - 1) / FTR) * ((ow + FS_TC - 1) / FS_TC);                  \
        if (!no_fixed && oh >= FTR && ow >= FS_TC && ftiles * batch >= 296) {                                     \
            P.tr = FTR; P.tc = FS_TC; P.pb = 1;                                                                   \
            geometry();                                                                                           \
            auto kern = wt_generic_synthesis_fixed_kernel<FL, MD, FTR>;                                           \
            WT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_FLOATS * 4));    \
            using G_ = FsGeom<FL, FTR>;                                                                           \
            CUtensorMap m0, m1, m2, m3;                                                                           \
            const bool t0 = make_plane_map(&m0, aa, batch, h2, w2, G_::CR, G_::CCS);                              \
            const bool t1 = make_plane_map(&m1, da, batch, h2, w2, G_::CR, G_::CCS);                              \
            const bool t2 = make_plane_map(&m2, ad, batch, h2, w2, G_::CR, G_::CCS);                              \
            const bool t3 = make_plane_map(&m3, dd, batch, h2, w2, G_::CR, G_::CCS);                              \
            P.use_tma = (t0 && t1 && t2 && t3) ? 1 : 0;                                                           \
            kern<<<(unsigned)(ftiles * batch), GEN_THREADS, (G_::SMEM_FLOATS_ + 4) * sizeof(float), stream>>>(bank, P, m0, m1, m2, m3); \
        } else {                                                                                                  \
            auto kern = wt_generic_synthesis_kernel<FL, MD>;                                                      \
            WT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_FLOATS * 4));    \
            kern<<<(unsigned)grid, GEN_THREADS, smem, stream>>>(bank, P);                                         \
        }                                                                                                         \
    }
    WT_GEN_DISPATCH(bank.F, mode)
#undef WT_GEN_LAUNCH
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

#ifndef WT_GENERIC_FMA
int wt_count_nonzero_launch(const float* x, long long n, long long* counts4, cudaStream_t stream) {
    if (n <= 0) return WT_OK;
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    long long blocks = std::min((n + 255) / 256, (long long)di.sm_count * 8);
    wt_count_nonzero_kernel<<<(int)blocks, 256, 0, stream>>>(x, n, counts4);
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}

// flat coefficient layout: [cA_L | L: da, ad, dd | ... | 1: da, ad, dd]
#endif  // !WT_GENERIC_FMA

static void band_offsets(const WtGenericLayer& L, int F, int mode, int* hs, int* ws, long long* off /*[level+1][3]*/,
                         long long* off_a) {
    wt_level_sizes(L.H, F, mode, L.level, hs);
    wt_level_sizes(L.W, F, mode, L.level, ws);
    long long o = 0;
    *off_a = 0;
    o += L.batch * hs[L.level] * ws[L.level];
    for (int l = L.level; l >= 1; --l) {
        for (int b = 0; b < 3; ++b) {
            off[l * 3 + b] = o;
            o += L.batch * hs[l] * ws[l];
        }
    }
}

int WT_GEN_FN(wt_generic_analysis)(const WtBank& bank, int mode, const WtGenericLayer& L, long long* counts, cudaStream_t stream,
                        int* launches) {
    int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1];
    long long off[(WT_MAX_LEVEL + 1) * 3], off_a;
    band_offsets(L, bank.F, mode, hs, ws, off, &off_a);
    int rc = WT_OK;
    if (L.level < 1) {
        rc = wt_count_nonzero_launch(L.in, L.batch * L.H * L.W, counts + (size_t)L.layer * 4, stream);
        if (rc) return rc;
        ++*launches;
    }
    WtView cur = wt_view(const_cast<float*>(L.in), L.H, L.W);
    for (int l = 1; l <= L.level; ++l) {
        float* dst = (l == L.level) ? L.coef + off_a : ((l & 1) ? L.tmp0 : L.tmp1);
        WtView aa = wt_view(dst, hs[l], ws[l]);
        WtView da = wt_view(L.coef + off[l * 3 + 0], hs[l], ws[l]);
        WtView ad = wt_view(L.coef + off[l * 3 + 1], hs[l], ws[l]);
        WtView dd = wt_view(L.coef + off[l * 3 + 2], hs[l], ws[l]);
        // the scratch of the two-sweep modes sits behind the ping-pong part of the buffer `cur` is NOT in
        float* scratch = wt_generic_scratch((l & 1) ? L.tmp0 : L.tmp1, L.batch, L.H, L.W, bank.F, mode, L.level);
        // level 1 reads the whole input anyway: the fixed-geometry kernel counts its nonzeros (counts[3], dwt_pruning_NoEntropy.py:58)
        // from the staged tiles; when another kernel takes the level, the count is a pass of its own
        int counted = 0;
        rc = WT_GEN_FN(wt_generic_analysis_level)(bank, mode, cur, L.batch, hs[l - 1], ws[l - 1], aa, da, ad, dd, scratch, stream,
                                       l == 1 ? counts + (size_t)L.layer * 4 : nullptr, l == 1 ? &counted : nullptr);
        if (rc) return rc;
        *launches += (mode > WT_MODE_SYMMETRIC) ? 2 : 1;
        if (l == 1 && !counted) {
            rc = wt_count_nonzero_launch(L.in, L.batch * L.H * L.W, counts + (size_t)L.layer * 4, stream);
            if (rc) return rc;
            ++*launches;
        }
        cur = aa;
    }
    return WT_OK;
}

int WT_GEN_FN(wt_generic_synthesis)(const WtBank& bank, int mode, const WtGenericLayer& L, const float* thresholds,
                         long long* counts, cudaStream_t stream, int* launches) {
    int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1];
    long long off[(WT_MAX_LEVEL + 1) * 3], off_a;
    band_offsets(L, bank.F, mode, hs, ws, off, &off_a);
    WtView cur = wt_view(L.coef + off_a, hs[L.level], ws[L.level]);
    const float* thr_dev = L.use_abs ? nullptr : thresholds + L.layer;
    for (int l = L.level; l >= 1; --l) {
        // reconstruction is cropped to the next level's detail size, or to the input shape at l == 1
        const int oh = hs[l - 1], ow = ws[l - 1];
        float* dst = (l == 1) ? L.out : ((l & 1) ? L.tmp0 : L.tmp1);
        WtView out = wt_view(dst, oh, ow);
        WtView da = wt_view(L.coef + off[l * 3 + 0], hs[l], ws[l]);
        WtView ad = wt_view(L.coef + off[l * 3 + 1], hs[l], ws[l]);
        WtView dd = wt_view(L.coef + off[l * 3 + 2], hs[l], ws[l]);
        int rc = WT_GEN_FN(wt_generic_synthesis_level)(bank, mode, cur, da, ad, dd, L.batch, hs[l], ws[l], out, oh, ow, thr_dev,
                                            L.abs_threshold, 1, l == L.level, l == 1 ? counts + (size_t)L.layer * 4 : nullptr,
                                            stream);
        if (rc) return rc;
        ++*launches;
        cur = out;
    }
    return WT_OK;
}
