/*
 * oracle/dwt_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, fp32, single thread) of the arithmetic the reference's DWT-prune hot
 * path executes:  pywt.wavedec2 -> pywt.coeffs_to_array -> np.percentile/np.where ->
 * pywt.array_to_coeffs -> pywt.waverec2, as called from
 *     /root/reference/ResNet/dwt_pruning.py:67-77       (percentile flow)
 *     /root/reference/ResNet/dwt_pruning_NoEntropy.py:41-49  (absolute-threshold flow)
 *
 * The arithmetic itself lives in PyWavelets 1.4.1 and NumPy 1.22.4 (requirements.txt:2-3), which
 * are NOT vendored under /root/reference and are NOT installed in this image.  This file restates
 * their published algorithms.  PARITY UNPINNED: the reference holds no
 * test, golden vector or fixture for this path, and real pywt cannot be run here; the oracle is
 * pinned only by perfect-reconstruction identities, the PyWavelets documentation examples
 * (tests/test_oracle.py) and the count laws of the reference's stored logs.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (wavelettransforms_b200/) never does.
 *
 * Conventions (definition of "bit-exact" for the CUDA path):
 *   - taps are (float) casts of the double tables; data and accumulators are float;
 *   - multiply then add, NO fused multiply-add (build with -ffp-contract=off);
 *   - analysis: one accumulator per output starting at 0.0f, taps in ascending index j;
 *   - periodization synthesis: every product is accumulated directly into out[o]
 *     (low-pass products in ascending j first, then high-pass products);
 *   - symmetric synthesis (and every other non-periodization mode: pywt's idwt does not look at
 *     the mode): a local sum per (output, filter) in ascending j, then out[o] += sum, low-pass
 *     before high-pass;
 *   - modes zero / constant / reflect / periodic / smooth / antisymmetric / antireflect: the
 *     extended sample is formed first (ext_sample: an index map, a negation, or one subtract +
 *     multiply + add for smooth, running subtract / add for antireflect), then multiplied by the tap;
 *   - dwt2: axis -2 first, then axis -1.  idwt2: axis -1 first, then axis -2.
 *
 * Arithmetic twin (orc_set_arithmetic(1), "fma_taporder"): the SAME tap order with every multiply-add
 * contracted into one fused multiply-add, acc = fmaf(tap, sample, acc) -- the oracle of the library's
 * named fast mode (wt_plan_set_arithmetic).  It is NOT the reference's arithmetic: exact zeros of a
 * round trip are not preserved; it is held to the faithful result by tolerance only.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "wavelet_tables.h"

#define ORC_MODE_PERIODIZATION 0
#define ORC_MODE_SYMMETRIC 1
/* the rest of pywt.Modes: signal extension as tabulated in the PyWavelets
 * documentation, "Signal extension modes"; all of them share the symmetric mode's lengths,
 * indexing and synthesis (pywt's idwt is the same for every mode but periodization) */
#define ORC_MODE_ZERO 2
#define ORC_MODE_CONSTANT 3
#define ORC_MODE_REFLECT 4
#define ORC_MODE_PERIODIC 5
#define ORC_MODE_SMOOTH 6
#define ORC_MODE_ANTISYMMETRIC 7
#define ORC_MODE_ANTIREFLECT 8
#define ORC_NUM_MODES 9
#define ORC_MAX_LEVEL 32

static int g_arith = 0; /* 0 faithful (multiply, then add), 1 fma_taporder */
void orc_set_arithmetic(int a) { g_arith = a ? 1 : 0; }
int orc_get_arithmetic(void) { return g_arith; }
/* acc + tap * v in the selected arithmetic */
static inline float mac(float acc, float tap, float v) {
    if (g_arith) return fmaf(tap, v, acc);
    float p = tap * v;
    return acc + p;
}

int orc_wavelet_index(const char* name) {
    for (int i = 0; i < ORC_NUM_WAVELETS; ++i)
        if (strcmp(name, ORC_BANKS[i].name) == 0) return i;
    return -1;
}

int orc_dec_len(int wid) {
    if (wid < 0 || wid >= ORC_NUM_WAVELETS) return -1;
    return ORC_BANKS[wid].len;
}

/* copies the four double filters; returns the length */
int orc_filters(int wid, double* dec_lo, double* dec_hi, double* rec_lo, double* rec_hi) {
    if (wid < 0 || wid >= ORC_NUM_WAVELETS) return -1;
    const orc_bank_t* b = &ORC_BANKS[wid];
    for (int i = 0; i < b->len; ++i) {
        dec_lo[i] = b->dec_lo[i]; dec_hi[i] = b->dec_hi[i];
        rec_lo[i] = b->rec_lo[i]; rec_hi[i] = b->rec_hi[i];
    }
    return b->len;
}

/* pywt.dwt_max_level: 0 if F<=1 or n<F-1 else floor(log2(n // (F-1))) */
int orc_dwt_max_level(int n, int F) {
    if (F <= 1 || n < F - 1) return 0;
    int q = n / (F - 1), l = 0;
    while (q > 1) { q >>= 1; ++l; }
    return l;
}

/* pywt.dwt_coeff_len */
int orc_dwt_coeff_len(int n, int F, int mode) {
    if (mode == ORC_MODE_PERIODIZATION) return (n + 1) / 2;
    return (n + F - 1) / 2;
}

/* pywt.idwt output length for coefficient length n */
int orc_idwt_len(int n, int F, int mode) {
    if (mode == ORC_MODE_PERIODIZATION) return 2 * n;
    return 2 * n - F + 2;
}

/* sizes[0] = n, sizes[l] = coefficient length at level l (l = 1..level) */
void orc_level_sizes(int n, int F, int mode, int level, int* sizes) {
    sizes[0] = n;
    for (int l = 1; l <= level; ++l) sizes[l] = orc_dwt_coeff_len(sizes[l - 1], F, mode);
}

/* side of pywt.coeffs_to_array's output along one transformed axis */
int orc_packed_side(int n, int F, int mode, int level) {
    if (level == 0) return n;
    int sizes[ORC_MAX_LEVEL + 1];
    orc_level_sizes(n, F, mode, level, sizes);
    int side = sizes[level];
    for (int l = level; l >= 1; --l) side += sizes[l];
    return side;
}

/* length along one axis of pywt.waverec2's output (before the reference's crop) */
int orc_rec_size(int n, int F, int mode, int level) {
    if (level == 0) return n;
    int sizes[ORC_MAX_LEVEL + 1];
    orc_level_sizes(n, F, mode, level, sizes);
    int a = sizes[level];
    for (int l = level; l >= 1; --l) {
        if (a == sizes[l] + 1) a = sizes[l]; /* waverec2 drops the extra approximation sample */
        if (a != sizes[l]) return -1;
        a = orc_idwt_len(a, F, mode);
    }
    return a;
}

static inline int pmod(int a, int m) { int r = a % m; return r < 0 ? r + m : r; }

/* whole-sample antisymmetric extension, the way pywt's convolution loops build it: a running edge
 * value `le` that each block of n-1 samples is mirrored around (first block around the edge sample
 * itself, the next one around the last extension value, ...).  d >= 1 = distance from the edge;
 * `near` / `far` = index of the edge sample and of the opposite end. */
static inline float antireflect_side(const float* x, int n, int xs, int d, int near, int far) {
    const float xn = x[(int64_t)near * xs], xf = x[(int64_t)far * xs];
    const int step = (far > near) ? 1 : -1;
    float le = xn;
    int blk = 0;
    while (d > n - 1) {                 /* whole blocks: le becomes the block's last value */
        float t = (blk & 1) ? (xn - xf) : (xf - xn);
        le = (blk & 1) ? (le + t) : (le - t);
        d -= n - 1;
        ++blk;
    }
    if (blk & 1) {                      /* odd blocks walk back from the far end */
        float t = x[(int64_t)(far - step * d) * xs] - xf;
        return le + t;
    } else {
        float t = x[(int64_t)(near + step * d) * xs] - xn;
        return le - t;
    }
}

/* value of the boundary-extended signal at (possibly out-of-range) index idx */
static inline float ext_sample(const float* x, int n, int xs, int idx, int mode) {
    if (mode == ORC_MODE_PERIODIZATION) {
        int np = n + (n & 1);              /* odd length: repeat the last sample once */
        int m = pmod(idx, np);
        if (m == n) m = n - 1;
        return x[(int64_t)m * xs];
    }
    if (idx >= 0 && idx < n) return x[(int64_t)idx * xs];
    switch (mode) {
    case ORC_MODE_ZERO:
        return 0.0f;
    case ORC_MODE_CONSTANT:
        return x[(int64_t)(idx < 0 ? 0 : n - 1) * xs];
    case ORC_MODE_REFLECT: {               /* whole-sample symmetric, period 2n-2 (n = 1: constant) */
        if (n == 1) return x[0];
        int per = 2 * n - 2, m = pmod(idx, per);
        if (m >= n) m = per - m;
        return x[(int64_t)m * xs];
    }
    case ORC_MODE_PERIODIC:
        return x[(int64_t)pmod(idx, n) * xs];
    case ORC_MODE_SMOOTH: {                /* straight line through the two edge samples (n = 1: constant) */
        if (n < 2) return x[0];
        if (idx < 0) {
            float slope = x[0] - x[xs];
            float t = slope * (float)(-idx);
            return x[0] + t;
        } else {
            const float e = x[(int64_t)(n - 1) * xs];
            float slope = e - x[(int64_t)(n - 2) * xs];
            float t = slope * (float)(idx - (n - 1));
            return e + t;
        }
    }
    case ORC_MODE_ANTISYMMETRIC: {         /* half-sample antisymmetric, period 2n */
        int m = pmod(idx, 2 * n);
        return m < n ? x[(int64_t)m * xs] : -x[(int64_t)(2 * n - 1 - m) * xs];
    }
    case ORC_MODE_ANTIREFLECT:             /* whole-sample antisymmetric (n = 1: constant) */
        if (n == 1) return x[0];
        return idx < 0 ? antireflect_side(x, n, xs, -idx, 0, n - 1)
                       : antireflect_side(x, n, xs, idx - (n - 1), n - 1, 0);
    default: {                             /* ORC_MODE_SYMMETRIC: half-sample symmetric, period 2n */
        int m = pmod(idx, 2 * n);
        if (m >= n) m = 2 * n - 1 - m;
        return x[(int64_t)m * xs];
    }
    }
}

/* pywt.pad(x, (before, after), mode) for the non-periodization modes: the extended signal the
 * analysis convolves (exposed so the tests can hold it against numpy.pad) */
int orc_extend(const float* x, int n, int mode, int before, int after, float* out) {
    if (n < 1 || mode < ORC_MODE_SYMMETRIC || mode >= ORC_NUM_MODES || before < 0 || after < 0) return -1;
    for (int i = -before; i < n + after; ++i) out[i + before] = ext_sample(x, n, 1, i, mode);
    return n + before + after;
}

/* one band of a single-level 1-D analysis */
static void dwt1d_band(const float* x, int n, int xs, const float* f, int F, int mode,
                       float* out, int os) {
    int no = orc_dwt_coeff_len(n, F, mode);
    int base = (mode == ORC_MODE_PERIODIZATION) ? F / 2 : 1;
    for (int o = 0; o < no; ++o) {
        float acc = 0.0f;
        for (int j = 0; j < F; ++j) acc = mac(acc, f[j], ext_sample(x, n, xs, base + 2 * o - j, mode));
        out[(int64_t)o * os] = acc;
    }
}

/* single-level 1-D synthesis; out must hold orc_idwt_len(n) floats with stride os */
static void idwt1d(const float* ca, const float* cd, int n, int cs, const float* rlo,
                   const float* rhi, int F, int mode, float* out, int os) {
    int no = orc_idwt_len(n, F, mode);
    for (int o = 0; o < no; ++o) out[(int64_t)o * os] = 0.0f;
    if (mode == ORC_MODE_PERIODIZATION) {
        int start = F / 4;
        int shift = ((F / 2) % 2 == 0) ? 1 : 0;
        for (int pass = 0; pass < 2; ++pass) {
            const float* c = pass ? cd : ca;
            const float* r = pass ? rhi : rlo;
            for (int i = start; i < n + start; ++i) {
                int oe = pmod(2 * (i - start) + shift, 2 * n);
                int oo = pmod(2 * (i - start) + shift + 1, 2 * n);
                for (int j = 0; j < F / 2; ++j) {
                    float v = c[(int64_t)pmod(i - j, n) * cs];
                    out[(int64_t)oe * os] = mac(out[(int64_t)oe * os], r[2 * j], v);
                    out[(int64_t)oo * os] = mac(out[(int64_t)oo * os], r[2 * j + 1], v);
                }
            }
        }
    } else {
        for (int pass = 0; pass < 2; ++pass) {
            const float* c = pass ? cd : ca;
            const float* r = pass ? rhi : rlo;
            for (int i = F / 2 - 1; i < n; ++i) {
                float se = 0.0f, so = 0.0f;
                for (int j = 0; j < F / 2; ++j) {
                    float v = c[(int64_t)(i - j) * cs];
                    se = mac(se, r[2 * j], v);
                    so = mac(so, r[2 * j + 1], v);
                }
                int o = 2 * (i - F / 2 + 1);
                out[(int64_t)o * os] = out[(int64_t)o * os] + se;
                out[(int64_t)(o + 1) * os] = out[(int64_t)(o + 1) * os] + so;
            }
        }
    }
}

typedef struct { float dlo[ORC_MAX_FILTER_LEN], dhi[ORC_MAX_FILTER_LEN], rlo[ORC_MAX_FILTER_LEN], rhi[ORC_MAX_FILTER_LEN]; int F; } bank32_t;

static int load_bank(int wid, bank32_t* b) {
    if (wid < 0 || wid >= ORC_NUM_WAVELETS) return -1;
    const orc_bank_t* s = &ORC_BANKS[wid];
    b->F = s->len;
    for (int i = 0; i < s->len; ++i) {
        b->dlo[i] = (float)s->dec_lo[i]; b->dhi[i] = (float)s->dec_hi[i];
        b->rlo[i] = (float)s->rec_lo[i]; b->rhi[i] = (float)s->rec_hi[i];
    }
    return 0;
}

/* Single-level 1-D transforms exposed for the known-answer tests (pywt.dwt / pywt.idwt). */
int orc_dwt(const float* x, int n, int wid, int mode, float* ca, float* cd) {
    bank32_t b; if (load_bank(wid, &b) || n < 1) return -1;
    dwt1d_band(x, n, 1, b.dlo, b.F, mode, ca, 1);
    dwt1d_band(x, n, 1, b.dhi, b.F, mode, cd, 1);
    return orc_dwt_coeff_len(n, b.F, mode);
}

int orc_idwt(const float* ca, const float* cd, int n, int wid, int mode, float* out) {
    bank32_t b; if (load_bank(wid, &b) || n < 1) return -1;
    if (mode != ORC_MODE_PERIODIZATION && n < b.F / 2) return -1;
    idwt1d(ca, cd, n, 1, b.rlo, b.rhi, b.F, mode, out, 1);
    return orc_idwt_len(n, b.F, mode);
}

/* dwt2 on one plane x[h][w] (row stride xw): bands aa, da, ad, dd each [h2][w2] contiguous.
 * tmp must hold 2*h2*w floats. */
static void dwt2_plane(const float* x, int h, int w, int xw, const bank32_t* b, int mode,
                       float* aa, float* da, float* ad, float* dd, float* tmp) {
    int h2 = orc_dwt_coeff_len(h, b->F, mode), w2 = orc_dwt_coeff_len(w, b->F, mode);
    float* lo = tmp;                 /* [h2][w] : 'a' along axis -2 */
    float* hi = tmp + (size_t)h2 * w; /* [h2][w] : 'd' along axis -2 */
    for (int c = 0; c < w; ++c) {
        dwt1d_band(x + c, h, xw, b->dlo, b->F, mode, lo + c, w);
        dwt1d_band(x + c, h, xw, b->dhi, b->F, mode, hi + c, w);
    }
    for (int r = 0; r < h2; ++r) {
        dwt1d_band(lo + (size_t)r * w, w, 1, b->dlo, b->F, mode, aa + (size_t)r * w2, 1);
        dwt1d_band(lo + (size_t)r * w, w, 1, b->dhi, b->F, mode, ad + (size_t)r * w2, 1);
        dwt1d_band(hi + (size_t)r * w, w, 1, b->dlo, b->F, mode, da + (size_t)r * w2, 1);
        dwt1d_band(hi + (size_t)r * w, w, 1, b->dhi, b->F, mode, dd + (size_t)r * w2, 1);
    }
}

/* idwt2 of four bands [h2][w2] (row strides given) -> out [H][W] contiguous with
 * H = idwt_len(h2), W = idwt_len(w2).  tmp holds 2*h2*W floats. */
static void idwt2_plane(const float* aa, int aas, const float* da, int das, const float* ad, int ads,
                        const float* dd, int dds, int h2, int w2, const bank32_t* b, int mode,
                        float* out, float* tmp) {
    int W = orc_idwt_len(w2, b->F, mode);
    float* lo = tmp;                    /* [h2][W] : 'a' along axis -2 */
    float* hi = tmp + (size_t)h2 * W;   /* [h2][W] : 'd' along axis -2 */
    for (int r = 0; r < h2; ++r) {      /* axis -1 first */
        idwt1d(aa + (size_t)r * aas, ad + (size_t)r * ads, w2, 1, b->rlo, b->rhi, b->F, mode, lo + (size_t)r * W, 1);
        idwt1d(da + (size_t)r * das, dd + (size_t)r * dds, w2, 1, b->rlo, b->rhi, b->F, mode, hi + (size_t)r * W, 1);
    }
    for (int c = 0; c < W; ++c)         /* then axis -2 */
        idwt1d(lo + c, hi + c, h2, W, b->rlo, b->rhi, b->F, mode, out + c, W);
}

/*
 * wavedec2 over the last two axes + coeffs_to_array(padding=0) (dwt_pruning.py:67-70).
 * x: [batch][H][W]; packed: [batch][SH][SW] with SH/SW = orc_packed_side(H/W, ...).
 * level 0 returns a copy of x (pywt returns the same array object).
 */
int orc_wavedec2_packed(const float* x, int64_t batch, int H, int W, int wid, int level, int mode,
                        float* packed) {
    bank32_t b;
    if (load_bank(wid, &b) || level < 0 || level > ORC_MAX_LEVEL || H < 1 || W < 1) return -1;
    if (level == 0) { memcpy(packed, x, (size_t)batch * H * W * sizeof(float)); return 0; }
    int hs[ORC_MAX_LEVEL + 1], ws[ORC_MAX_LEVEL + 1];
    orc_level_sizes(H, b.F, mode, level, hs);
    orc_level_sizes(W, b.F, mode, level, ws);
    int SH = orc_packed_side(H, b.F, mode, level), SW = orc_packed_side(W, b.F, mode, level);
    int mh = H, mw = W;
    for (int l = 1; l <= level; ++l) { if (hs[l] > mh) mh = hs[l]; if (ws[l] > mw) mw = ws[l]; }
    size_t maxp = (size_t)mh * mw;
    size_t maxt = 2 * maxp;
    float* cur = (float*)malloc(maxp * sizeof(float));
    float* nxt = (float*)malloc(maxp * sizeof(float));
    float* bands = (float*)malloc(3 * maxp * sizeof(float));
    float* tmp = (float*)malloc(maxt * sizeof(float));
    memset(packed, 0, (size_t)batch * SH * SW * sizeof(float));
    /* offsets of the detail blocks: level L sits right after cA_L, level 1 is outermost */
    int offh[ORC_MAX_LEVEL + 1], offw[ORC_MAX_LEVEL + 1];
    { int ah = hs[level], aw = ws[level];
      for (int l = level; l >= 1; --l) { offh[l] = ah; offw[l] = aw; ah += hs[l]; aw += ws[l]; } }
    for (int64_t p = 0; p < batch; ++p) {
        const float* src = x + (size_t)p * H * W;
        float* dst = packed + (size_t)p * SH * SW;
        int ch = H, cw = W; const float* in = src;
        for (int l = 1; l <= level; ++l) {
            int h2 = hs[l], w2 = ws[l];
            float* da = bands; float* ad = bands + (size_t)h2 * w2; float* dd = bands + 2 * (size_t)h2 * w2;
            dwt2_plane(in, ch, cw, cw, &b, mode, nxt, da, ad, dd, tmp);
            for (int r = 0; r < h2; ++r) {
                memcpy(dst + (size_t)(offh[l] + r) * SW, da + (size_t)r * w2, w2 * sizeof(float));
                memcpy(dst + (size_t)r * SW + offw[l], ad + (size_t)r * w2, w2 * sizeof(float));
                memcpy(dst + (size_t)(offh[l] + r) * SW + offw[l], dd + (size_t)r * w2, w2 * sizeof(float));
            }
            float* t = cur; cur = nxt; nxt = t; in = cur; ch = h2; cw = w2;
        }
        for (int r = 0; r < ch; ++r) memcpy(dst + (size_t)r * SW, cur + (size_t)r * cw, cw * sizeof(float));
    }
    free(cur); free(nxt); free(bands); free(tmp);
    return 0;
}

/*
 * array_to_coeffs + waverec2 (dwt_pruning.py:75-77).  packed: [batch][SH][SW];
 * out: [batch][RH][RW] with RH/RW = orc_rec_size(H/W, ...)  (the caller crops, as the reference does).
 */
int orc_waverec2_packed(const float* packed, int64_t batch, int H, int W, int wid, int level, int mode,
                        float* out) {
    bank32_t b;
    if (load_bank(wid, &b) || level < 0 || level > ORC_MAX_LEVEL || H < 1 || W < 1) return -1;
    if (level == 0) { memcpy(out, packed, (size_t)batch * H * W * sizeof(float)); return 0; }
    int hs[ORC_MAX_LEVEL + 1], ws[ORC_MAX_LEVEL + 1];
    orc_level_sizes(H, b.F, mode, level, hs);
    orc_level_sizes(W, b.F, mode, level, ws);
    int SH = orc_packed_side(H, b.F, mode, level), SW = orc_packed_side(W, b.F, mode, level);
    int RH = orc_rec_size(H, b.F, mode, level), RW = orc_rec_size(W, b.F, mode, level);
    if (RH < 0 || RW < 0) return -2;
    int mh = H + 1, mw = W + 1;
    for (int l = 1; l <= level; ++l) {
        int ih = orc_idwt_len(hs[l], b.F, mode), iw = orc_idwt_len(ws[l], b.F, mode);
        if (hs[l] > mh) mh = hs[l];
        if (ws[l] > mw) mw = ws[l];
        if (ih > mh) mh = ih;
        if (iw > mw) mw = iw;
    }
    size_t maxp = (size_t)mh * mw;
    float* cur = (float*)malloc(maxp * sizeof(float));
    float* nxt = (float*)malloc(maxp * sizeof(float));
    float* tmp = (float*)malloc(2 * maxp * sizeof(float));
    int offh[ORC_MAX_LEVEL + 1], offw[ORC_MAX_LEVEL + 1];
    { int ah = hs[level], aw = ws[level];
      for (int l = level; l >= 1; --l) { offh[l] = ah; offw[l] = aw; ah += hs[l]; aw += ws[l]; } }
    for (int64_t p = 0; p < batch; ++p) {
        const float* src = packed + (size_t)p * SH * SW;
        int ah = hs[level], aw = ws[level];  /* current approximation size and row stride */
        int astride = aw;
        for (int r = 0; r < ah; ++r) memcpy(cur + (size_t)r * aw, src + (size_t)r * SW, aw * sizeof(float));
        for (int l = level; l >= 1; --l) {
            int h2 = hs[l], w2 = ws[l];
            /* waverec2: drop the extra trailing approximation sample when a is d+1;
             * done by reading only h2 x w2 of cur with its original row stride. */
            (void)ah; (void)aw;
            const float* da = src + (size_t)offh[l] * SW;
            const float* ad = src + offw[l];
            const float* dd = src + (size_t)offh[l] * SW + offw[l];
            idwt2_plane(cur, astride, da, SW, ad, SW, dd, SW, h2, w2, &b, mode, nxt, tmp);
            float* t = cur; cur = nxt; nxt = t;
            ah = orc_idwt_len(h2, b.F, mode); aw = orc_idwt_len(w2, b.F, mode); astride = aw;
        }
        memcpy(out + (size_t)p * RH * RW, cur, (size_t)RH * RW * sizeof(float));
    }
    free(cur); free(nxt); free(tmp);
    return 0;
}

/*
 * np.percentile(|c|, p) for a float32 population, given the two order statistics.
 * semantics 0 = legacy numpy 1.22 (the reference's pin): q, virtual index, gamma and the lerp in
 * float64 (b-a rounded to float32 first); the caller's `<` then demotes t to float32.
 * semantics 1 = NEP 50 / numpy 2.x: everything float32.
 * Returns the threshold already rounded to the float32 the comparison uses.
 */
void orc_percentile_plan(int64_t n, double p, int semantics, int64_t* k0, int64_t* k1, double* gamma) {
    if (semantics == 0) {
        double q = p / 100.0;
        double vi = (double)(n - 1) * q;
        if (vi >= (double)(n - 1)) { *k0 = n - 1; *k1 = n - 1; *gamma = vi - floor(vi); return; }
        double f = floor(vi);
        *k0 = (int64_t)f; *k1 = *k0 + 1; *gamma = vi - f;
    } else {
        float q = (float)p / 100.0f;
        float nm1 = (float)(n - 1);
        float vi = nm1 * q;
        float f = floorf(vi);
        float g = vi - f;
        *gamma = (double)g;
        if (vi >= nm1) { *k0 = n - 1; *k1 = n - 1; return; }
        *k0 = (int64_t)f; *k1 = *k0 + 1;
        if (*k1 > n - 1) *k1 = n - 1;
    }
}

float orc_percentile_lerp(float a, float b, double gamma, int semantics) {
    if (semantics == 0) {
        float d32 = b - a;
        double d = (double)d32, t;
        if (gamma >= 0.5) { double m = d * (1.0 - gamma); t = (double)b - m; }
        else { double m = d * gamma; t = (double)a + m; }
        return (float)t;
    } else {
        float g = (float)gamma, d = b - a, t;
        if (g >= 0.5f) { float m = d * (1.0f - g); t = b - m; }
        else { float m = d * g; t = a + m; }
        return t;
    }
}

static int cmp_float(const void* x, const void* y) {
    float a = *(const float*)x, b = *(const float*)y;
    return (a > b) - (a < b);
}

/* full flow of percentile_based_thresholding (dwt_pruning.py:25-32) on a float32 array, in place:
 * returns the float32 threshold used by the strict `<` compare. */
float orc_percentile_threshold(float* c, int64_t n, double p, int semantics) {
    float* a = (float*)malloc((size_t)n * sizeof(float));
    for (int64_t i = 0; i < n; ++i) a[i] = fabsf(c[i]);
    qsort(a, (size_t)n, sizeof(float), cmp_float);
    int64_t k0, k1; double gamma;
    orc_percentile_plan(n, p, semantics, &k0, &k1, &gamma);
    float t = orc_percentile_lerp(a[k0], a[k1], gamma, semantics);
    free(a);
    for (int64_t i = 0; i < n; ++i) if (fabsf(c[i]) < t) c[i] = 0.0f;
    return t;
}

/* absolute threshold of dwt_pruning_NoEntropy.py:45, in place */
void orc_absolute_threshold(float* c, int64_t n, float thr) {
    for (int64_t i = 0; i < n; ++i) if (fabsf(c[i]) < thr) c[i] = 0.0f;
}
