// wt_small_core.cuh -- register-resident multi-level 2-D DWT / IDWT of one small plane (kH x kW conv-kernel planes,
// SURVEY F4), shared by the fused prune kernels (wt_fused.cuh) and the stand-alone analysis / synthesis kernels.
//
// Replaces pywt.wavedec2 / pywt.waverec2 over axes (-2,-1) of a conv weight [O,I,kH,kW]
// (/root/reference/ResNet/dwt_pruning.py:67-68,77; dwt_pruning_NoEntropy.py:41,49), mode 'periodization', including
// N < F where every tap wraps several times (F6).  One thread owns one plane; every loop bound is a template constant, so
// after unrolling every array index is static and every filter tap is a constant-bank operand.
//
// Arithmetic follows oracle/dwt_oracle.c exactly (fp32, multiply then add, no FMA, same order) in the default
// `Faithful` arithmetic; `FmaTapOrder` (plan option arithmetic = fma_taporder) contracts every multiply-add into one
// fused multiply-add IN THE SAME TAP ORDER (oracle twin: dwt_oracle.c built with WT_ORACLE_FMA, std::fmaf).
#pragma once
#include "wt_common.cuh"

namespace wtsmall {
// ------------------------------------------------------------------------------------------------
// compile-time index helpers
// ------------------------------------------------------------------------------------------------
__host__ __device__ constexpr int cpmod(int a, int m) { return ((a % m) + m) % m; }
__host__ __device__ constexpr int half_up(int n) { return (n + 1) / 2; }
// index of the periodized (odd length padded by repeating the last sample) signal
__host__ __device__ constexpr int per_index(int idx, int n) {
    int np = n + (n & 1);
    int m = cpmod(idx, np);
    return m == n ? n - 1 : m;
}

// Compile-time zero-tap masks of a filter bank (bit j set = tap j is exactly 0.0f).  A product with a
// zero tap is +-0 and the accumulators start at +0.0f and can never become -0.0f (x + (-x) = +0 under
// round-to-nearest), so skipping those terms leaves every result bit-identical for finite inputs; the
// launcher derives the masks from the actual taps, so a specialised instantiation is only used when they match.
// FMA = true selects the `fma_taporder` arithmetic: acc = fmaf(tap, v, acc) in the same tap order (one rounding per
// term instead of two; exact zeros of the faithful arithmetic are not preserved).
template <unsigned DL, unsigned DH, unsigned RL, unsigned RH, bool FMA = false>
struct ZeroTaps {
    static constexpr unsigned dl = DL, dh = DH, rl = RL, rh = RH;
    static constexpr bool fma = FMA;
};
using NoZeroTaps = ZeroTaps<0, 0, 0, 0>;
using NoZeroTapsFma = ZeroTaps<0, 0, 0, 0, true>;
#define WT_TAP(mask, j) ((((mask) >> (j)) & 1u) == 0u)

// Packed fp32x2 additions (Blackwell FADD2): two independent round-to-nearest fp32 additions in one issue slot.
// Only the ADDS are packed: ptxas contracts mul.rn.f32x2 feeding add.rn.f32x2 into FFMA2 even under -fmad=false,
// which would change the last bit; scalar FMULs feeding an FADD2 stay as written (checked in the SASS).  The two
// halves are ordinary registers, so a lane of a pair can still be updated alone by a scalar FADD.
__device__ __forceinline__ void fadd2(float& a0, float& a1, float p0, float p1) {
    const float2 r = __fadd2_rn(make_float2(a0, a1), make_float2(p0, p1));
    a0 = r.x;
    a1 = r.y;
}
// acc += tap * v in the arithmetic Z selects
template <class Z>
__device__ __forceinline__ float mac(float acc, float tap, float v) {
    if constexpr (Z::fma) return __fmaf_rn(tap, v, acc);
    else return __fadd_rn(acc, __fmul_rn(tap, v));
}
// (a0, a1) += (t0 * v0, t1 * v1): packed adds (faithful) or one packed FFMA2 (fma_taporder)
template <class Z>
__device__ __forceinline__ void mac2(float& a0, float& a1, float t0, float v0, float t1, float v1) {
    if constexpr (Z::fma) {
        const float2 r = __ffma2_rn(make_float2(t0, t1), make_float2(v0, v1), make_float2(a0, a1));
        a0 = r.x;
        a1 = r.y;
    } else {
        fadd2(a0, a1, __fmul_rn(t0, v0), __fmul_rn(t1, v1));
    }
}

// single-level 2-D analysis of an H x W plane held in registers (dwt2: axis -2 first, then -1)
template <int H, int W, int F, class Z>
__device__ __forceinline__ void dwt2_regs(const float (&x)[H * W], const WtBank& bk, float (&aa)[half_up(H) * half_up(W)],
                                          float (&da)[half_up(H) * half_up(W)], float (&ad)[half_up(H) * half_up(W)],
                                          float (&dd)[half_up(H) * half_up(W)]) {
    constexpr int H2 = half_up(H), W2 = half_up(W);
    float lo[H2 * W], hi[H2 * W];
#pragma unroll
    for (int c = 0; c < W; ++c) {
#pragma unroll
        for (int o = 0; o < H2; ++o) {
            float sl = 0.0f, sh = 0.0f;
#pragma unroll
            for (int j = 0; j < F; ++j) {
                const float v = x[per_index(F / 2 + 2 * o - j, H) * W + c];
                if (WT_TAP(Z::dl, j) && WT_TAP(Z::dh, j)) {
                    mac2<Z>(sl, sh, bk.dec_lo[j], v, bk.dec_hi[j], v);
                } else {
                    if (WT_TAP(Z::dl, j)) sl = mac<Z>(sl, bk.dec_lo[j], v);
                    if (WT_TAP(Z::dh, j)) sh = mac<Z>(sh, bk.dec_hi[j], v);
                }
            }
            lo[o * W + c] = sl;
            hi[o * W + c] = sh;
        }
    }
#pragma unroll
    for (int r = 0; r < H2; ++r) {
#pragma unroll
        for (int o = 0; o < W2; ++o) {
            float saa = 0.0f, sad = 0.0f, sda = 0.0f, sdd = 0.0f;
#pragma unroll
            for (int j = 0; j < F; ++j) {
                const int m = per_index(F / 2 + 2 * o - j, W);
                const float vl = lo[r * W + m], vh = hi[r * W + m];
                if (WT_TAP(Z::dl, j)) mac2<Z>(saa, sda, bk.dec_lo[j], vl, bk.dec_lo[j], vh);
                if (WT_TAP(Z::dh, j)) mac2<Z>(sad, sdd, bk.dec_hi[j], vl, bk.dec_hi[j], vh);
            }
            aa[r * W2 + o] = saa;
            ad[r * W2 + o] = sad;
            da[r * W2 + o] = sda;
            dd[r * W2 + o] = sdd;
        }
    }
}

// one output sample of the periodization synthesis of (a, d), coefficient length N, taken from
// register arrays with element stride S starting at offset OFF (direct accumulation)
template <int N, int F, int O, int S, int OFF, int LEN, class Z>
__device__ __forceinline__ float idwt_per_sample(const float (&a)[LEN], const float (&d)[LEN], const WtBank& bk) {
    constexpr int start = F / 4;
    constexpr int shift = ((F / 2) % 2 == 0) ? 1 : 0;
    constexpr int q = cpmod(O - shift, 2 * N);
    constexpr int i = start + q / 2;
    constexpr int par = q & 1;
    float acc = 0.0f;
#pragma unroll
    for (int j = 0; j < F / 2; ++j)
        if (WT_TAP(Z::rl, 2 * j + par)) acc = mac<Z>(acc, bk.rec_lo[2 * j + par], a[OFF + cpmod(i - j, N) * S]);
#pragma unroll
    for (int j = 0; j < F / 2; ++j)
        if (WT_TAP(Z::rh, 2 * j + par)) acc = mac<Z>(acc, bk.rec_hi[2 * j + par], d[OFF + cpmod(i - j, N) * S]);
    return acc;
}

template <int H, int W, int F, int R, int C, class Z>
struct RowSynth {  // axis -1 pass: lo[r][c], hi[r][c] for r < H2, c < W (row-major, W columns kept)
    static constexpr int H2 = half_up(H), W2 = half_up(W);
    __device__ __forceinline__ static void run(const float (&aa)[H2 * W2], const float (&da)[H2 * W2],
                                               const float (&ad)[H2 * W2], const float (&dd)[H2 * W2], const WtBank& bk,
                                               float (&lo)[H2 * W], float (&hi)[H2 * W]) {
        lo[R * W + C] = idwt_per_sample<W2, F, C, 1, R * W2, H2 * W2, Z>(aa, ad, bk);
        hi[R * W + C] = idwt_per_sample<W2, F, C, 1, R * W2, H2 * W2, Z>(da, dd, bk);
        if constexpr (C + 1 < W) RowSynth<H, W, F, R, C + 1, Z>::run(aa, da, ad, dd, bk, lo, hi);
        else if constexpr (R + 1 < H2) RowSynth<H, W, F, R + 1, 0, Z>::run(aa, da, ad, dd, bk, lo, hi);
    }
};

template <int H, int W, int F, int R, int C, class Z>
struct ColSynth {  // axis -2 pass: x[r][c] for r < H, c < W
    static constexpr int H2 = half_up(H);
    __device__ __forceinline__ static void run(const float (&lo)[H2 * W], const float (&hi)[H2 * W], const WtBank& bk,
                                               float (&x)[H * W]) {
        x[R * W + C] = idwt_per_sample<H2, F, R, W, C, H2 * W, Z>(lo, hi, bk);
        if constexpr (C + 1 < W) ColSynth<H, W, F, R, C + 1, Z>::run(lo, hi, bk, x);
        else if constexpr (R + 1 < H) ColSynth<H, W, F, R + 1, 0, Z>::run(lo, hi, bk, x);
    }
};

// single-level 2-D synthesis cropped to H x W (idwt2: axis -1 first, then axis -2)
template <int H, int W, int F, class Z>
__device__ __forceinline__ void idwt2_regs(const float (&aa)[half_up(H) * half_up(W)], const float (&da)[half_up(H) * half_up(W)],
                                           const float (&ad)[half_up(H) * half_up(W)], const float (&dd)[half_up(H) * half_up(W)],
                                           const WtBank& bk, float (&x)[H * W]) {
    constexpr int H2 = half_up(H);
    float lo[H2 * W], hi[H2 * W];
    RowSynth<H, W, F, 0, 0, Z>::run(aa, da, ad, dd, bk, lo, hi);
    ColSynth<H, W, F, 0, 0, Z>::run(lo, hi, bk, x);
}

// ------------------------------------------------------------------------------------------------
// multi-level recursion.  Coefficient numbering: level-1 details (da, ad, dd) first, then level 2,
// ..., finally cA_L; all offsets are compile-time at the depth where the recursion stops.
// ------------------------------------------------------------------------------------------------
template <int H, int W, int F, int DEPTH, int BASE, class Z>
struct SmallPlane {
    static constexpr int H2 = half_up(H), W2 = half_up(W), NB = H2 * W2;

    template <class Put>
    __device__ __forceinline__ static void analyze(const float (&x)[H * W], int level, const WtBank& bk, Put& put) {
        if (level == 0) {
#pragma unroll
            for (int i = 0; i < H * W; ++i) put(BASE + i, x[i]);
            return;
        }
        if constexpr (DEPTH > 0) {
            float aa[NB], da[NB], ad[NB], dd[NB];
            dwt2_regs<H, W, F, Z>(x, bk, aa, da, ad, dd);
#pragma unroll
            for (int i = 0; i < NB; ++i) {
                put(BASE + i, da[i]);
                put(BASE + NB + i, ad[i]);
                put(BASE + 2 * NB + i, dd[i]);
            }
            SmallPlane<H2, W2, F, DEPTH - 1, BASE + 3 * NB, Z>::analyze(aa, level - 1, bk, put);
        }
    }

    template <class Get>
    __device__ __forceinline__ static void synthesize(float (&x)[H * W], int level, const WtBank& bk, Get& get) {
        if (level == 0) {
#pragma unroll
            for (int i = 0; i < H * W; ++i) x[i] = get(BASE + i);
            return;
        }
        if constexpr (DEPTH > 0) {
            float aa[NB], da[NB], ad[NB], dd[NB];
#pragma unroll
            for (int i = 0; i < NB; ++i) {
                da[i] = get(BASE + i);
                ad[i] = get(BASE + NB + i);
                dd[i] = get(BASE + 2 * NB + i);
            }
            // the deeper reconstruction is already cropped to H2 x W2: that is waverec2's
            // "drop the extra approximation sample" rule
            SmallPlane<H2, W2, F, DEPTH - 1, BASE + 3 * NB, Z>::synthesize(aa, level - 1, bk, get);
            idwt2_regs<H, W, F, Z>(aa, da, ad, dd, bk, x);
        }
    }
};

}  // namespace wtsmall
