// wt_generic.cuh -- generic (any plane size, every pywt mode, any level) DWT path.
#pragma once
#include "wt_common.cuh"

// strided view of a batch of planes: element (p, r, c) lives at p[pl*pstride + r*rstride + c]
struct WtView {
    float* p;
    long long pstride;
    int rstride;
};

static inline WtView wt_view(float* p, int rows, int cols) {
    WtView v;
    v.p = p;
    v.pstride = (long long)rows * cols;
    v.rstride = cols;
    return v;
}

struct WtGenericLayer {
    const float* in;
    float* out;
    long long batch;
    int H, W, level, layer;
    float* coef;  // [cA_L | level L: da, ad, dd | ... | level 1: da, ad, dd], each band [batch][h_l][w_l]
    float* tmp0;  // ping-pong approximation buffers, wt_generic_tmp_elems() floats each
    float* tmp1;
    float abs_threshold;
    int use_abs;
};

// per-plane number of coefficients / tmp elements
long long wt_generic_ncoef(int H, int W, int F, int mode, int level);
long long wt_generic_tmp_elems(int H, int W, int F, int mode, int level);
// modes beyond symmetric: where the scratch of the two-sweep analysis starts inside a ping-pong buffer of
// batch * wt_generic_tmp_elems() floats (nullptr for periodization / symmetric)
float* wt_generic_scratch(float* tmp, long long batch, int H, int W, int F, int mode, int level);

// one decomposition level: in (h x w) -> aa, da, ad, dd (h2 x w2); scratch: see wt_generic_scratch
// nz_counts4 / nz_counted (optional): when the kernel that takes the level can count the nonzeros of `in` on the way it adds
// them to nz_counts4[3], sets nz_counts4[0] = numel and *nz_counted = 1; otherwise *nz_counted = 0 and the caller counts
int wt_generic_analysis_level(const WtBank& bank, int mode, WtView in, long long batch, int h, int w, WtView aa,
                              WtView da, WtView ad, WtView dd, float* scratch, cudaStream_t stream,
                              long long* nz_counts4 = nullptr, int* nz_counted = nullptr);
// one reconstruction level: bands (h2 x w2) -> out (oh x ow, cropped).  thr: device threshold
// (nullable) or thr_host when use_host; thr_a: also threshold the approximation band (deepest
// level only); counts4 != nullptr: add zero / nonzero counts of `out` (final level).
int wt_generic_synthesis_level(const WtBank& bank, int mode, WtView aa, WtView da, WtView ad, WtView dd,
                               long long batch, int h2, int w2, WtView out, int oh, int ow, const float* thr_dev,
                               float thr_host, int use_thr, int thr_a, long long* counts4, cudaStream_t stream);

// whole-layer drivers over the flat coefficient layout; *launches += kernels enqueued
int wt_generic_analysis(const WtBank& bank, int mode, const WtGenericLayer& L, long long* counts, cudaStream_t stream,
                        int* launches);
int wt_generic_synthesis(const WtBank& bank, int mode, const WtGenericLayer& L, const float* thresholds,
                         long long* counts, cudaStream_t stream, int* launches);

// the same four in the named fast mode fma_taporder (wt_generic_fma.cu: one FMA per tap, same tap order)
int wt_generic_analysis_level_fma(const WtBank& bank, int mode, WtView in, long long batch, int h, int w, WtView aa,
                                  WtView da, WtView ad, WtView dd, float* scratch, cudaStream_t stream,
                                  long long* nz_counts4 = nullptr, int* nz_counted = nullptr);
int wt_generic_synthesis_level_fma(const WtBank& bank, int mode, WtView aa, WtView da, WtView ad, WtView dd,
                                   long long batch, int h2, int w2, WtView out, int oh, int ow, const float* thr_dev,
                                   float thr_host, int use_thr, int thr_a, long long* counts4, cudaStream_t stream);
int wt_generic_analysis_fma(const WtBank& bank, int mode, const WtGenericLayer& L, long long* counts, cudaStream_t stream,
                            int* launches);
int wt_generic_synthesis_fma(const WtBank& bank, int mode, const WtGenericLayer& L, const float* thresholds,
                             long long* counts, cudaStream_t stream, int* launches);

// counts4[3] += count(x != 0); counts4[0] = n
int wt_count_nonzero_launch(const float* x, long long n, long long* counts4, cudaStream_t stream);
