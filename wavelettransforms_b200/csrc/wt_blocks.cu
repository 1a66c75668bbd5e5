// wt_blocks.cu -- building blocks exposed through the C ABI for tests and for callers that want the
// coefficient array itself:
//   wt_select_abs_f32        np.partition inside np.percentile (dwt_pruning.py:27)
//   wt_wavedec2_packed_f32   pywt.wavedec2 + pywt.coeffs_to_array (dwt_pruning.py:67-70)
//   wt_waverec2_packed_f32   pywt.array_to_coeffs + pywt.waverec2 + crop (dwt_pruning.py:75-82)
// They allocate their scratch with cudaMalloc and synchronise the stream before freeing it, so they
// are convenience entry points, not the hot path (plans are).
#include <algorithm>

#include "wt_common.cuh"
#include "wt_generic.cuh"

extern "C" int wt_select_abs_f32(const float* x, int64_t n, int64_t implicit_zeros, int64_t k0, int64_t k1, float* ab_out,
                                 void* stream_) {
    WT_CHECK(n >= 0 && implicit_zeros >= 0 && n + implicit_zeros >= 1, WT_E_INVALID, "empty population");
    WT_CHECK(ab_out != nullptr && (n == 0 || x != nullptr), WT_E_INVALID, "NULL argument");
    WT_CHECK(k0 >= 0 && k1 >= k0 && k1 < n + implicit_zeros, WT_E_INVALID, "ranks out of range");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n == 0) {
        WT_CUDA(cudaMemsetAsync(ab_out, 0, 2 * sizeof(float), stream));
        return WT_OK;
    }
    const size_t scr = (wt_select_scratch_bytes(1) + 255) / 256 * 256;
    const size_t o_job = 0, o_thr = 256, o_cnt = 512, o_scr = 1024;
    const unsigned int cap = wt_select_cand_cap(n);
    const size_t o_cand = o_scr + scr;
    char* dev = nullptr;
    WT_CUDA(cudaMalloc((void**)&dev, o_cand + (size_t)cap * sizeof(unsigned int)));
    WtSelJob j;
    memset(&j, 0, sizeof(j));
    j.x = x; j.out = nullptr; j.n = n; j.zeros = implicit_zeros; j.k0 = k0; j.k1 = k1; j.gamma = 0.0;
    j.lerp_mode = WT_LERP_NEP50; j.tile_begin = 0;
    j.cand = reinterpret_cast<unsigned int*>(dev + o_cand);
    j.cand_cap = cap;
    j.sample_shift = wt_select_sample_shift(n);
    const long long nt = (n + WT_SEL_TILE - 1) / WT_SEL_TILE;
    int rc = WT_OK;
    if (nt >= 2147483647LL) { wt_set_error("too many tiles"); rc = WT_E_UNSUPPORTED; }
    j.n_tiles = (int)nt; j.layer = 0;
    cudaError_t e = cudaSuccess;
    if (!rc) {
        WtSelArgs a;
        memset(&a, 0, sizeof(a));
        if ((e = cudaMemcpyAsync(dev + o_job, &j, sizeof(j), cudaMemcpyHostToDevice, stream)) == cudaSuccess &&
            (e = cudaMemsetAsync(dev + o_thr, 0, o_scr - o_thr + scr, stream)) == cudaSuccess) {
            wt_select_bind(&a, dev + o_scr, 1);
            a.jobs = reinterpret_cast<WtSelJob*>(dev + o_job);
            a.n_jobs = 1; a.n_tiles = (int)nt; a.any_apply = 0;
            a.thresholds = reinterpret_cast<float*>(dev + o_thr);
            a.counts = reinterpret_cast<long long*>(dev + o_cnt);
            rc = wt_select_launch(a, stream);
            if (!rc) e = cudaMemcpyAsync(ab_out, &a.state->a, 2 * sizeof(float), cudaMemcpyDeviceToDevice, stream);
        }
    }
    cudaError_t e2 = cudaStreamSynchronize(stream);
    cudaFree(dev);
    if (!rc && (e != cudaSuccess || e2 != cudaSuccess)) {
        wt_set_error("wt_select_abs_f32: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
        rc = WT_E_CUDA;
    }
    return rc;
}

namespace {
struct PackedGeom {
    int F, level, mode;
    int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1], offh[WT_MAX_LEVEL + 1], offw[WT_MAX_LEVEL + 1];
    int SH, SW;
};

int packed_geom(int H, int W, int wid, int level, int mode, PackedGeom* g, WtBank* bank) {
    int rc = wt_get_bank(wid, bank);
    if (rc) return rc;
    WT_CHECK(mode >= 0 && mode < WT_NUM_MODES, WT_E_INVALID, "Unknown mode name.");
    WT_CHECK(level >= 0, WT_E_INVALID, "Level value of %d is too low . Minimum level is 0.", level);
    WT_CHECK(level <= WT_MAX_LEVEL, WT_E_UNSUPPORTED, "level %d above the supported maximum %d", level, WT_MAX_LEVEL);
    WT_CHECK(H >= 1 && W >= 1, WT_E_INVALID, "empty plane");
    g->F = bank->F; g->level = level; g->mode = mode;
    wt_level_sizes(H, g->F, mode, level, g->hs);
    wt_level_sizes(W, g->F, mode, level, g->ws);
    int ah = g->hs[level], aw = g->ws[level];
    for (int l = level; l >= 1; --l) {
        g->offh[l] = ah; g->offw[l] = aw;
        ah += g->hs[l]; aw += g->ws[l];
    }
    g->SH = level ? ah : H;
    g->SW = level ? aw : W;
    return WT_OK;
}

WtView sub_view(float* base, int SH, int SW, int r0, int c0) {
    WtView v;
    v.p = base + (long long)r0 * SW + c0;
    v.pstride = (long long)SH * SW;
    v.rstride = SW;
    return v;
}
}  // namespace

extern "C" int wt_wavedec2_packed_f32(const float* x, int64_t batch, int H, int W, int wid, int level, int mode,
                                      float* packed, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    PackedGeom g;
    WtBank bank;
    int rc = packed_geom(H, W, wid, level, mode, &g, &bank);
    if (rc) return rc;
    WT_CHECK(batch >= 0, WT_E_INVALID, "negative batch");
    if (batch == 0) return WT_OK;
    WT_CHECK(x && packed, WT_E_INVALID, "NULL argument");
    if (level == 0) {
        WT_CUDA(cudaMemcpyAsync(packed, x, (size_t)batch * H * W * sizeof(float), cudaMemcpyDeviceToDevice, stream));
        return WT_OK;
    }
    const long long tmp_elems = batch * wt_generic_tmp_elems(H, W, g.F, mode, level);
    float* tmp = nullptr;
    WT_CUDA(cudaMalloc((void**)&tmp, (size_t)tmp_elems * 2 * sizeof(float)));
    cudaError_t e = cudaMemsetAsync(packed, 0, (size_t)batch * g.SH * g.SW * sizeof(float), stream);
    WtView cur = wt_view(const_cast<float*>(x), H, W);
    for (int l = 1; l <= level && !rc && e == cudaSuccess; ++l) {
        WtView aa = (l == level) ? sub_view(packed, g.SH, g.SW, 0, 0) : wt_view(tmp + (l & 1) * tmp_elems, g.hs[l], g.ws[l]);
        WtView da = sub_view(packed, g.SH, g.SW, g.offh[l], 0);
        WtView ad = sub_view(packed, g.SH, g.SW, 0, g.offw[l]);
        WtView dd = sub_view(packed, g.SH, g.SW, g.offh[l], g.offw[l]);
        float* scratch = wt_generic_scratch(tmp + (l & 1) * tmp_elems, batch, H, W, g.F, mode, level);
        rc = wt_generic_analysis_level(bank, mode, cur, batch, g.hs[l - 1], g.ws[l - 1], aa, da, ad, dd, scratch, stream);
        cur = aa;
    }
    cudaError_t e2 = cudaStreamSynchronize(stream);
    cudaFree(tmp);
    if (!rc && (e != cudaSuccess || e2 != cudaSuccess)) {
        wt_set_error("wt_wavedec2_packed_f32: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
        rc = WT_E_CUDA;
    }
    return rc;
}

extern "C" int wt_waverec2_packed_f32(const float* packed, int64_t batch, int H, int W, int wid, int level, int mode,
                                      float* out, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    PackedGeom g;
    WtBank bank;
    int rc = packed_geom(H, W, wid, level, mode, &g, &bank);
    if (rc) return rc;
    WT_CHECK(batch >= 0, WT_E_INVALID, "negative batch");
    if (batch == 0) return WT_OK;
    WT_CHECK(out && packed, WT_E_INVALID, "NULL argument");
    if (level == 0) {
        WT_CUDA(cudaMemcpyAsync(out, packed, (size_t)batch * H * W * sizeof(float), cudaMemcpyDeviceToDevice, stream));
        return WT_OK;
    }
    const long long tmp_elems = batch * wt_generic_tmp_elems(H, W, g.F, mode, level);
    float* tmp = nullptr;
    WT_CUDA(cudaMalloc((void**)&tmp, (size_t)tmp_elems * 2 * sizeof(float)));
    float* pk = const_cast<float*>(packed);
    WtView cur = sub_view(pk, g.SH, g.SW, 0, 0);
    for (int l = level; l >= 1 && !rc; --l) {
        const int oh = g.hs[l - 1], ow = g.ws[l - 1];
        WtView o = (l == 1) ? wt_view(out, oh, ow) : wt_view(tmp + (l & 1) * tmp_elems, oh, ow);
        WtView da = sub_view(pk, g.SH, g.SW, g.offh[l], 0);
        WtView ad = sub_view(pk, g.SH, g.SW, 0, g.offw[l]);
        WtView dd = sub_view(pk, g.SH, g.SW, g.offh[l], g.offw[l]);
        rc = wt_generic_synthesis_level(bank, mode, cur, da, ad, dd, batch, g.hs[l], g.ws[l], o, oh, ow, nullptr, 0.0f, 0, 0,
                                        nullptr, stream);
        cur = o;
    }
    cudaError_t e2 = cudaStreamSynchronize(stream);
    cudaFree(tmp);
    if (!rc && e2 != cudaSuccess) {
        wt_set_error("wt_waverec2_packed_f32: %s", cudaGetErrorString(e2));
        rc = WT_E_CUDA;
    }
    return rc;
}

// ---- wt_key_hist_f32: one radix pass of a DISTRIBUTED order-statistic select --------------------------------------
// A tensor whose planes are split over several GPUs (intra-layer split) needs np.percentile over the union
// of every rank's coefficients.  Each rank histograms one digit of its own keys (bit pattern of |x|, 31 bits) under a
// prefix, the caller sums the histograms over the ranks (NCCL all-reduce of digit_bits <= 12 -> at most 32 KB) and walks
// down: three passes (12 + 10 + 9 bits) give the exact k-th smallest of the union.  hist is ACCUMULATED into (the caller
// zeroes it), so several tensors can share one histogram.
namespace {
constexpr int KH_THREADS = 512;
__global__ void __launch_bounds__(KH_THREADS) wt_key_hist_kernel(const float* __restrict__ x, long long n, unsigned int prefix,
                                                                 int prefix_bits, int digit_bits, unsigned long long* hist) {
    __shared__ unsigned int sh[4096];
    const int nb = 1 << digit_bits;
    for (int b = threadIdx.x; b < nb; b += KH_THREADS) sh[b] = 0u;
    __syncthreads();
    const int below = 31 - prefix_bits - digit_bits;          // key bits under the digit
    const unsigned int mask = (unsigned int)nb - 1u;
    const long long stride = (long long)gridDim.x * KH_THREADS;
    for (long long i = (long long)blockIdx.x * KH_THREADS + threadIdx.x; i < n; i += stride) {
        const unsigned int key = __float_as_uint(__ldcg(&x[i])) & 0x7fffffffu;
        if (prefix_bits == 0 || (key >> (31 - prefix_bits)) == prefix) atomicAdd(&sh[(key >> below) & mask], 1u);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += KH_THREADS) {
        const unsigned int v = sh[b];
        if (v) atomicAdd(&hist[b], (unsigned long long)v);
    }
}
}  // namespace

extern "C" int wt_key_hist_f32(const float* x, int64_t n, uint32_t prefix, int prefix_bits, int digit_bits, uint64_t* hist,
                               void* stream_) {
    WT_CHECK(n >= 0 && hist != nullptr && (n == 0 || x != nullptr), WT_E_INVALID, "bad argument");
    WT_CHECK(prefix_bits >= 0 && digit_bits >= 1 && digit_bits <= 12 && prefix_bits + digit_bits <= 31, WT_E_INVALID,
             "wt_key_hist_f32: prefix_bits + digit_bits must stay within the 31 key bits, digit_bits within 1..12");
    WT_CHECK(prefix_bits == 0 || prefix_bits == 31 || (prefix >> prefix_bits) == 0u, WT_E_INVALID, "prefix wider than prefix_bits");
    if (n == 0) return WT_OK;
    WtDeviceInfo di;
    int rc = wt_device_info(&di);
    if (rc) return rc;
    const long long want = (n + KH_THREADS * 8 - 1) / (KH_THREADS * 8);
    const int grid = (int)std::max<long long>(1, std::min<long long>(want, (long long)di.sm_count * 4));
    wt_key_hist_kernel<<<grid, KH_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(x, (long long)n, prefix, prefix_bits, digit_bits,
                                                                                     reinterpret_cast<unsigned long long*>(hist));
    WT_CUDA(cudaGetLastError());
    return WT_OK;
}
