// wt_api.cu -- C ABI of libwtprune: metadata, planning and the plan runner.
//
// A plan is the host-side mirror of one call of the reference's
//   multi_resolution_analysis(weights, wavelet, level, percentile|threshold, mode)
// (/root/reference/ResNet/dwt_pruning.py:35-95, dwt_pruning_NoEntropy.py:12-62) over a whole group
// of tensors: it classifies every tensor (level-0 select-and-zero, small-plane register path,
// generic tiled path), lays out one device block (job tables, histograms, coefficient workspace,
// optional host-I/O staging) and replays a fixed launch sequence with no host synchronisation.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>

#include "wt_common.cuh"
#include "wt_generic.cuh"
#include "wt_wavelet_tables.h"

// ---- errors ---------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

void wt_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* wt_last_error(void) { return g_err; }
extern "C" int wt_version(void) { return 1; }

int wt_device_info(WtDeviceInfo* info) {
    static thread_local int cached_dev = -1;
    static thread_local WtDeviceInfo cached;
    int dev = 0;
    WT_CUDA(cudaGetDevice(&dev));
    if (dev != cached_dev) {
        cached.device = dev;
        WT_CUDA(cudaDeviceGetAttribute(&cached.sm_count, cudaDevAttrMultiProcessorCount, dev));
        WT_CUDA(cudaDeviceGetAttribute(&cached.coop, cudaDevAttrCooperativeLaunch, dev));
        cached_dev = dev;
    }
    *info = cached;
    return WT_OK;
}

// ---- wavelets -----------------------------------------------------------------------------------------
extern "C" int wt_wavelet_lookup(const char* name, int* id, int* dec_len) {
    WT_CHECK(name != nullptr, WT_E_INVALID, "wavelet name is NULL");
    for (int i = 0; i < WT_NUM_WAVELETS; ++i) {
        if (strcmp(name, WT_BANKS[i].name) == 0) {
            if (id) *id = i;
            if (dec_len) *dec_len = WT_BANKS[i].len;
            return WT_OK;
        }
    }
    wt_set_error("Unknown wavelet name '%s', check wavelist() for the list of available builtin wavelets.", name);
    return WT_E_INVALID;
}

extern "C" int wt_wavelet_filters(int wid, double* dec_lo, double* dec_hi, double* rec_lo, double* rec_hi) {
    WT_CHECK(wid >= 0 && wid < WT_NUM_WAVELETS, WT_E_INVALID, "bad wavelet id %d", wid);
    const wt_bank_t& b = WT_BANKS[wid];
    for (int i = 0; i < b.len; ++i) {
        if (dec_lo) dec_lo[i] = b.dec_lo[i];
        if (dec_hi) dec_hi[i] = b.dec_hi[i];
        if (rec_lo) rec_lo[i] = b.rec_lo[i];
        if (rec_hi) rec_hi[i] = b.rec_hi[i];
    }
    return WT_OK;
}

int wt_get_bank(int wid, WtBank* bank) {
    WT_CHECK(wid >= 0 && wid < WT_NUM_WAVELETS, WT_E_INVALID, "bad wavelet id %d", wid);
    const wt_bank_t& b = WT_BANKS[wid];
    memset(bank, 0, sizeof(*bank));
    bank->F = b.len;
    for (int i = 0; i < b.len; ++i) {
        bank->dec_lo[i] = (float)b.dec_lo[i];
        bank->dec_hi[i] = (float)b.dec_hi[i];
        bank->rec_lo[i] = (float)b.rec_lo[i];
        bank->rec_hi[i] = (float)b.rec_hi[i];
    }
    return WT_OK;
}

extern "C" int wt_dwt_max_level(int n, int F) {
    if (F <= 1 || n < F - 1) return 0;
    int q = n / (F - 1), l = 0;
    while (q > 1) {
        q >>= 1;
        ++l;
    }
    return l;
}

extern "C" int wt_dwt_coeff_len(int n, int F, int mode) {
    return mode == WT_MODE_PERIODIZATION ? (n + 1) / 2 : (n + F - 1) / 2;
}

int wt_idwt_len(int n, int F, int mode) { return mode == WT_MODE_PERIODIZATION ? 2 * n : 2 * n - F + 2; }

void wt_level_sizes(int n, int F, int mode, int level, int* sizes) {
    sizes[0] = n;
    for (int l = 1; l <= level; ++l) sizes[l] = wt_dwt_coeff_len(sizes[l - 1], F, mode);
}

extern "C" int wt_packed_shape(int H, int W, int wid, int level, int mode, int* side_h, int* side_w,
                               int64_t* coeffs_per_plane) {
    WT_CHECK(wid >= 0 && wid < WT_NUM_WAVELETS, WT_E_INVALID, "bad wavelet id %d", wid);
    WT_CHECK(mode >= 0 && mode < WT_NUM_MODES, WT_E_INVALID, "Unknown mode name.");
    WT_CHECK(level >= 0, WT_E_INVALID, "Level value of %d is too low . Minimum level is 0.", level);
    WT_CHECK(level <= WT_MAX_LEVEL, WT_E_UNSUPPORTED, "level %d above the supported maximum %d", level, WT_MAX_LEVEL);
    WT_CHECK(H >= 1 && W >= 1, WT_E_INVALID, "empty plane");
    const int F = WT_BANKS[wid].len;
    int hs[WT_MAX_LEVEL + 1], ws[WT_MAX_LEVEL + 1];
    wt_level_sizes(H, F, mode, level, hs);
    wt_level_sizes(W, F, mode, level, ws);
    int sh = hs[level], sw = ws[level];
    int64_t nc = (int64_t)hs[level] * ws[level];
    for (int l = level; l >= 1; --l) {
        sh += hs[l];
        sw += ws[l];
        nc += 3 * (int64_t)hs[l] * ws[l];
    }
    if (side_h) *side_h = sh;
    if (side_w) *side_w = sw;
    if (coeffs_per_plane) *coeffs_per_plane = nc;
    return WT_OK;
}

extern "C" int wt_percentile_plan(int64_t n, double p, int sem, int64_t* k0, int64_t* k1, double* gamma) {
    WT_CHECK(n >= 1, WT_E_INVALID, "percentile of an empty population");
    WT_CHECK(p >= 0.0 && p <= 100.0, WT_E_INVALID, "Percentiles must be in the range [0, 100]");
    if (sem == WT_NUMPY_LEGACY_1_22) {
        // numpy 1.22: q = p / 100 (float64), virtual index (n-1)*q in float64
        double q = p / 100.0;
        double vi = (double)(n - 1) * q;
        double f = floor(vi);
        *gamma = vi - f;
        if (vi >= (double)(n - 1)) {
            *k0 = *k1 = n - 1;
        } else {
            *k0 = (int64_t)f;
            *k1 = *k0 + 1;
        }
    } else if (sem == WT_NUMPY_NEP50_2X) {
        // numpy 2.x on a float32 array: q = float32(p) / float32(100), index arithmetic in float32
        volatile float q = (float)p / 100.0f;
        volatile float nm1 = (float)(n - 1);
        volatile float vi = nm1 * q;
        float f = floorf(vi);
        volatile float g = vi - f;
        *gamma = (double)g;
        if (vi >= nm1) {
            *k0 = *k1 = n - 1;
        } else {
            *k0 = (int64_t)f;
            *k1 = *k0 + 1;
            if (*k1 > n - 1) *k1 = n - 1;
        }
    } else {
        wt_set_error("unknown numpy_semantics %d", sem);
        return WT_E_INVALID;
    }
    return WT_OK;
}

// ---- plans ------------------------------------------------------------------------------------------------
enum { PATH_EMPTY = 0, PATH_LEVEL0 = 1, PATH_SMALL = 2, PATH_GENERIC = 3, PATH_FUSED = 4 };

// one persistent fused launch (wt_fused.cuh): the flat class (H = W = 0: every level-0 tensor) or one small-plane size
struct FusedClass {
    int H = 0, W = 0;
    std::vector<int> layers;       // plan layer of every job, in job order (ascending population)
    std::vector<WtFJob> jobs;
    WtFJob* jobs_dev = nullptr;
    char* scratch[2] = {nullptr, nullptr};
    size_t scratch_bytes = 0;
    int parity = 0;                // scratch block the next run uses (the run zeroes the other one for the run after)
    int run_parity = 0;            // ... and the one the run being enqueued uses
    int total_items = 0, total_s_items = 0, n_select = 0;
    double share = 1.0;            // this class's part of the bytes of all fused classes of the plan
    unsigned long long* stamps_dev = nullptr;  // WT_FUSED_STAMPS tuning aid
};

struct SmallClass {
    int H, W;
    std::vector<WtSmallJob> jobs;
    WtSmallJob* jobs_dev = nullptr;
    int n_tiles = 0;
};

struct WtPlan {
    std::vector<WtLayerDesc> layers;
    std::vector<int> path;
    int wavelet_id = 0, mode = 0, thr_kind = 0, sem = 0, host_io = 0;
    double thr_value = 0;
    WtBank bank;
    int device = 0;

    char* dev = nullptr;  // one allocation
    size_t dev_bytes = 0;

    // select
    std::vector<WtSelJob> sel_jobs;
    WtSelJob* sel_jobs_dev = nullptr;
    WtSelArgs sel_args;
    // zeroed every run: [select scratch | counts]
    WtPeers peers = {};  // layer-sharded runs: byte distance to every peer's mapping of the output buffer
    char* zero_block = nullptr;    // the block the last / current run uses
    size_t zero_bytes = 0;
    char* zero_blocks[2] = {nullptr, nullptr};  // percentile plans: two blocks alternate, the select kernel of run n zeroes run n + 1's
    size_t counts_off = 0;         // offset of the counts inside a zero block
    int zpar = 0;
    float* thresholds_dev = nullptr;
    long long* counts_dev = nullptr;

    std::vector<SmallClass> small;
    std::vector<FusedClass> fused;
    bool precompact = false;  // percentile plans with small-plane layers: the analysis kernels do the select's streaming work
    int arithmetic = 0;  // 0: faithful (no FMA, the oracle's arithmetic), 1: fma_taporder (wt_plan_set_arithmetic)
    bool legacy_zero = false;  // some launch still accumulates into the zeroed block (generic path, legacy kernels)
    std::vector<WtGenericLayer> generic;
    std::vector<int> level0_abs;  // layers pruned by the plain threshold kernel (absolute, level 0)

    // per-layer device pointers actually used by the kernels (staging for host I/O)
    std::vector<float*> d_in, d_out, d_coef;
    std::vector<unsigned int*> d_cand;  // candidate-key lists of the select fast path
    std::vector<unsigned int> cand_cap;
    std::vector<long long> ncoef, zeros;  // per layer: explicit coefficients / implicit padding zeros
    // generic layers are dealt round-robin to `gen_lanes` streams; a lane owns one pair of ping-pong buffers
    std::vector<float*> tmp0, tmp1;
    int gen_lanes = 1;
    int n_launches = 0;

    // optional per-phase device timing (wt_plan_set_timing): events[i] is recorded before phase i,
    // events.back() after the last one
    unsigned long long* stamps_dev = nullptr;  // WT_SELECT_STAMPS tuning aid
    int* cta_table_dev = nullptr;              // cost-balanced tile split of the select kernel: [grid + 1] + [grid]
    // the small-plane size classes (1x1, 3x3, 7x7 ...) are independent launches: they run concurrently on
    // auxiliary streams forked from / joined to the caller's stream by events
    std::vector<cudaStream_t> aux;
    std::vector<cudaEvent_t> ev_join;
    cudaEvent_t ev_fork = nullptr;

    // host-I/O plans are a pipeline over groups of layers: child g is a device-I/O plan over the staging
    // copies of its layers; H2D of group g+1, the kernels of group g and D2H of group g-1 overlap
    std::vector<WtPlan*> children;
    std::vector<std::vector<int>> child_layers;  // parent layer index of every child layer
    cudaStream_t s_in = nullptr, s_out = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_k;
    cudaEvent_t ev_start = nullptr, ev_done = nullptr;
    int timing = 0;
    // wt_plan_set_graph: the launch sequence of a run, captured once per scratch parity and replayed
    int use_graph = -1;            // -1: automatic (plans with large-plane layers, from the second run on), 0: never, 1: always
    unsigned long long run_count = 0;
    cudaGraphExec_t graph_exec[2] = {nullptr, nullptr};
    int graph_launches[2] = {0, 0};
    int graph_failed = 0;          // capture refused once (legacy default stream, ...): enqueue directly from then on
    std::vector<cudaEvent_t> events;
    std::vector<std::string> phase_names;
    std::vector<int> phase_launches;
    size_t n_phases = 0;
};

static size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

static long long layer_numel(const WtLayerDesc& L) { return L.batch * (long long)L.H * L.W; }

// job tables and ticket queues of the fused classes (wt_fused.cuh)
static int build_fused_tables(WtPlan* P) {
    const bool pct = (P->thr_kind == WT_THRESHOLD_PERCENTILE);
    for (auto& c : P->fused) {
        c.jobs.clear();
        c.n_select = 0;
        for (int i : c.layers) {
            const WtLayerDesc& L = P->layers[i];
            WtFJob j;
            memset(&j, 0, sizeof(j));
            j.in = P->d_in[i]; j.out = P->d_out[i];
            j.level = L.level; j.layer = i;
            j.abs_threshold = (float)P->thr_value;
            long long keys_per_unit = 1;
            if (c.H == 0) {
                j.units = layer_numel(L);
            } else {
                j.units = L.batch;
                keys_per_unit = P->ncoef[i] / L.batch;
                j.zeros = P->zeros[i];
            }
            wt_fused_geometry(c.H, c.W, &j, keys_per_unit);
            WT_CHECK(j.n < (1LL << 31), WT_E_UNSUPPORTED, "layer %d: population too large for the fused path", i);
            if (pct) {
                int64_t k0, k1; double gamma;
                int rc = wt_percentile_plan(j.n + j.zeros, P->thr_value, P->sem, &k0, &k1, &gamma);
                if (rc) return rc;
                j.k0 = k0; j.k1 = k1; j.gamma = gamma;
                j.lerp_mode = (P->sem == WT_NUMPY_LEGACY_1_22) ? WT_LERP_LEGACY : WT_LERP_NEP50;
                j.cand = P->d_cand[i];
                j.cand_cap = P->cand_cap[i];
                ++c.n_select;
            } else {
                j.lerp_mode = WT_LERP_ABSOLUTE;
            }
            j.peers = P->peers;
            c.jobs.push_back(j);
        }
        // the class-wide item spaces of the full passes and of the sample pass: every CTA takes one contiguous share
        c.total_items = c.total_s_items = 0;
        for (WtFJob& j : c.jobs) {
            j.item_begin = c.total_items;
            j.s_begin = c.total_s_items;
            c.total_items += j.items;
            c.total_s_items += j.s_items;
        }
        if (!c.jobs.empty())
            WT_CHECK(c.jobs.size() <= 4096, WT_E_UNSUPPORTED, "more than 4096 tensors in one fused class");
    }
    return WT_OK;
}

// fills the host-side job tables from plan->layers and the current device pointers
static int build_tables(WtPlan* P) {
    const int n = (int)P->layers.size();
    P->sel_jobs.clear();
    for (auto& c : P->small) { c.jobs.clear(); c.n_tiles = 0; }
    P->generic.clear();
    P->level0_abs.clear();
    int sel_tiles = 0;
    bool any_apply = false;
    {
        int rc = build_fused_tables(P);
        if (rc) return rc;
    }
    for (int i = 0; i < n; ++i) {
        const WtLayerDesc& L = P->layers[i];
        const long long numel = layer_numel(L);
        if (P->path[i] == PATH_EMPTY || P->path[i] == PATH_FUSED) continue;
        const bool pct = (P->thr_kind == WT_THRESHOLD_PERCENTILE);
        if (P->path[i] == PATH_SMALL) {
            SmallClass* cls = nullptr;
            for (auto& c : P->small) if (c.H == L.H && c.W == L.W) cls = &c;
            WtSmallJob j;
            memset(&j, 0, sizeof(j));
            j.in = P->d_in[i]; j.out = P->d_out[i]; j.coef = P->d_coef[i];
            j.batch = L.batch; j.level = L.level; j.layer = i;
            j.tile_begin = cls->n_tiles;
            j.n_tiles = (int)((L.batch + WT_SMALL_THREADS - 1) / WT_SMALL_THREADS);
            j.abs_threshold = (float)P->thr_value;
            j.use_abs = pct ? 0 : 1;
            j.peers = P->peers;
            j.sel_job = pct ? (int)P->sel_jobs.size() : -1;   // the select job of this tensor is pushed below, in the same iteration
            j.sample_shift = wt_select_sample_shift(P->ncoef[i]);
            j.cand = P->d_cand[i];
            j.cand_cap = P->cand_cap[i];
            cls->n_tiles += j.n_tiles;
            cls->jobs.push_back(j);
        } else if (P->path[i] == PATH_GENERIC) {
            WtGenericLayer g;
            memset(&g, 0, sizeof(g));
            g.in = P->d_in[i]; g.out = P->d_out[i]; g.batch = L.batch; g.H = L.H; g.W = L.W; g.level = L.level;
            g.layer = i; g.coef = P->d_coef[i];
            g.tmp0 = P->tmp0[P->generic.size() % P->gen_lanes]; g.tmp1 = P->tmp1[P->generic.size() % P->gen_lanes];
            g.abs_threshold = (float)P->thr_value; g.use_abs = pct ? 0 : 1;
            P->generic.push_back(g);
        } else if (!pct) {
            P->level0_abs.push_back(i);
        }
        if (pct) {
            WtSelJob j;
            memset(&j, 0, sizeof(j));
            const bool l0 = (P->path[i] == PATH_LEVEL0);
            j.x = l0 ? P->d_in[i] : P->d_coef[i];
            j.out = l0 ? P->d_out[i] : nullptr;
            j.n = l0 ? numel : P->ncoef[i];
            j.zeros = l0 ? 0 : P->zeros[i];
            int64_t k0, k1; double gamma;
            int rc = wt_percentile_plan(j.n + j.zeros, P->thr_value, P->sem, &k0, &k1, &gamma);
            if (rc) return rc;
            j.k0 = k0; j.k1 = k1; j.gamma = gamma;
            j.lerp_mode = (P->sem == WT_NUMPY_LEGACY_1_22) ? WT_LERP_LEGACY : WT_LERP_NEP50;
            j.cand = P->d_cand[i];
            j.cand_cap = P->cand_cap[i];
            j.sample_shift = wt_select_sample_shift(j.n);
            if (P->precompact && P->path[i] != PATH_SMALL) j.sample_shift = -1;   // not covered by the analysis kernels: radix passes
            if (l0) j.peers = P->peers;
            j.tile_begin = sel_tiles;
            const long long nt = (j.n + WT_SEL_TILE - 1) / WT_SEL_TILE;
            WT_CHECK(sel_tiles + nt < 2147483647LL, WT_E_UNSUPPORTED, "too many select tiles");
            j.n_tiles = (int)nt;
            j.layer = i;
            sel_tiles += (int)nt;
            any_apply |= l0;
            P->sel_jobs.push_back(j);
        }
    }
    P->sel_args.n_tiles = sel_tiles;
    P->sel_args.any_apply = any_apply ? 1 : 0;
    P->sel_args.short_lists = wt_select_predict_short(P->sel_jobs.data(), (int)P->sel_jobs.size());
    P->sel_args.precompacted = P->precompact ? 1 : 0;
    return WT_OK;
}

static int upload_tables(WtPlan* P) {
    if (!P->sel_jobs.empty()) {
        WT_CUDA(cudaMemcpy(P->sel_jobs_dev, P->sel_jobs.data(), P->sel_jobs.size() * sizeof(WtSelJob), cudaMemcpyHostToDevice));
        // cost-balanced split of the select's tile space (segments are not free: see WT_SEL_SEGMENT_COST)
        int grid = 1;
        int rc = wt_select_grid(P->sel_args.n_tiles, &grid);
        if (rc) return rc;
        std::vector<int> table((size_t)2 * grid + 1);
        wt_select_balance(P->sel_jobs.data(), (int)P->sel_jobs.size(), P->sel_args.n_tiles, grid, table.data(), table.data() + grid + 1);
        // the table must cover every tile exactly once and name the job of each CTA's first tile
        bool ok = table[0] == 0 && table[grid] == P->sel_args.n_tiles;
        for (int c = 0; ok && c < grid; ++c) {
            ok = table[c] <= table[c + 1];
            if (ok && table[c] < table[c + 1]) {
                const WtSelJob& jb = P->sel_jobs[table[grid + 1 + c]];
                ok = table[c] >= jb.tile_begin && table[c] < jb.tile_begin + jb.n_tiles;
            }
        }
        WT_CHECK(ok, WT_E_CUDA, "internal error: inconsistent select tile split");
        if (P->cta_table_dev) { cudaFree(P->cta_table_dev); P->cta_table_dev = nullptr; }
        WT_CUDA(cudaMalloc((void**)&P->cta_table_dev, table.size() * sizeof(int)));
        WT_CUDA(cudaMemcpy(P->cta_table_dev, table.data(), table.size() * sizeof(int), cudaMemcpyHostToDevice));
        P->sel_args.cta_tiles = P->cta_table_dev;
        P->sel_args.cta_job = P->cta_table_dev + grid + 1;
        P->sel_args.cta_grid = grid;
    }
    for (auto& c : P->small)
        if (!c.jobs.empty())
            WT_CUDA(cudaMemcpy(c.jobs_dev, c.jobs.data(), c.jobs.size() * sizeof(WtSmallJob), cudaMemcpyHostToDevice));
    for (auto& c : P->fused) {
        if (c.jobs.empty()) continue;
        WT_CUDA(cudaMemcpy(c.jobs_dev, c.jobs.data(), c.jobs.size() * sizeof(WtFJob), cudaMemcpyHostToDevice));
    }
    return WT_OK;
}

static void drop_graphs(WtPlan* P);
#define WT_GRAPH_AUTO_MIN_LAUNCHES 4

extern "C" int wt_plan_destroy(WtPlan* P) {
    if (!P) return WT_OK;
    for (WtPlan* c : P->children) wt_plan_destroy(c);
    for (cudaEvent_t e : P->ev_in) cudaEventDestroy(e);
    for (cudaEvent_t e : P->ev_k) cudaEventDestroy(e);
    if (P->ev_start) cudaEventDestroy(P->ev_start);
    if (P->ev_done) cudaEventDestroy(P->ev_done);
    if (P->s_in) cudaStreamDestroy(P->s_in);
    if (P->s_out) cudaStreamDestroy(P->s_out);
    for (int g = 0; g < 2; ++g) if (P->graph_exec[g]) cudaGraphExecDestroy(P->graph_exec[g]);
    for (cudaEvent_t e : P->events) cudaEventDestroy(e);
    for (cudaEvent_t e : P->ev_join) cudaEventDestroy(e);
    if (P->ev_fork) cudaEventDestroy(P->ev_fork);
    for (cudaStream_t q : P->aux) cudaStreamDestroy(q);
    for (auto& c : P->fused) if (c.stamps_dev) cudaFree(c.stamps_dev);
    if (P->stamps_dev) cudaFree(P->stamps_dev);
    if (P->cta_table_dev) cudaFree(P->cta_table_dev);
    if (P->dev) cudaFree(P->dev);
    delete P;
    return WT_OK;
}

// ---- host-I/O plans: a copy/compute pipeline over groups of layers ----------------------------------------
#define WT_PIPE_MAX_GROUPS 16

static int create_host_pipeline(const WtLayerDesc* layers, int n_layers, int wavelet_id, int mode, int threshold_kind,
                                double threshold_value, int numpy_semantics, WtPlan** plan_out) {
    WtPlan* P = new WtPlan();
    int rc = wt_get_bank(wavelet_id, &P->bank);
    if (rc) { delete P; return rc; }
    if (cudaGetDevice(&P->device) != cudaSuccess) {
        delete P;
        wt_set_error("no usable CUDA device (libwtprune has no CPU fallback): %s", cudaGetErrorString(cudaGetLastError()));
        return WT_E_CUDA;
    }
    P->wavelet_id = wavelet_id; P->mode = mode; P->thr_kind = threshold_kind; P->thr_value = threshold_value;
    P->sem = numpy_semantics; P->host_io = 1;
    P->layers.assign(layers, layers + n_layers);
    P->path.assign(n_layers, PATH_EMPTY);
    P->d_in.assign(n_layers, nullptr);
    size_t bytes = 0, total = 0;
    std::vector<size_t> off(n_layers, 0);
    std::vector<int> live;
    for (int i = 0; i < n_layers; ++i) {
        const WtLayerDesc& L = P->layers[i];
        if (L.batch < 0 || L.H < 0 || L.W < 0) { delete P; wt_set_error("layer %d: negative shape", i); return WT_E_INVALID; }
        const long long numel = layer_numel(L);
        if (numel == 0) continue;
        if (L.in == nullptr || L.out == nullptr) { delete P; wt_set_error("layer %d: NULL tensor pointer", i); return WT_E_INVALID; }
        P->path[i] = PATH_LEVEL0;  // "not empty"; the children classify for real
        off[i] = bytes;
        bytes = align_up(bytes + (size_t)numel * sizeof(float));
        total += (size_t)numel * sizeof(float);
        live.push_back(i);
    }
    P->dev_bytes = bytes;
    cudaError_t e = cudaMalloc((void**)&P->dev, std::max(bytes, (size_t)256));
    if (e != cudaSuccess) {
        wt_set_error("cudaMalloc(%zu) failed: %s (libwtprune has no CPU fallback)", bytes, cudaGetErrorString(e));
        delete P;
        return WT_E_CUDA;
    }
    for (int i : live) P->d_in[i] = reinterpret_cast<float*>(P->dev + off[i]);
    // contiguous groups of about total / G bytes, in layer order
    int gmax = WT_PIPE_MAX_GROUPS;
    if (const char* env = getenv("WT_PIPE_GROUPS")) gmax = std::max(1, atoi(env));
    const int G = std::max(1, std::min<int>(gmax, (int)live.size()));
    const size_t target = (total + G - 1) / G;
    std::vector<std::vector<int>> groups;
    size_t acc = 0;
    for (int i : live) {
        if (groups.empty() || (acc >= target && (int)groups.size() < G)) { groups.emplace_back(); acc = 0; }
        groups.back().push_back(i);
        acc += (size_t)layer_numel(P->layers[i]) * sizeof(float);
    }
    rc = WT_OK;
    for (auto& g : groups) {
        std::vector<WtLayerDesc> d;
        for (int i : g) {
            WtLayerDesc c = P->layers[i];
            c.in = P->d_in[i];
            c.out = P->d_in[i];  // pruned in place inside the staging copy
            d.push_back(c);
        }
        WtPlan* child = nullptr;
        rc = wt_plan_create(d.data(), (int)d.size(), wavelet_id, mode, threshold_kind, threshold_value, numpy_semantics, 0, &child);
        if (rc) break;
        P->children.push_back(child);
        P->child_layers.push_back(g);
    }
    if (!rc) {
        bool ok = cudaStreamCreateWithFlags(&P->s_in, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&P->s_out, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaEventCreateWithFlags(&P->ev_start, cudaEventDisableTiming) == cudaSuccess &&
                  cudaEventCreateWithFlags(&P->ev_done, cudaEventDisableTiming) == cudaSuccess;
        for (size_t g = 0; ok && g < groups.size(); ++g) {
            cudaEvent_t a, b;
            ok = cudaEventCreateWithFlags(&a, cudaEventDisableTiming) == cudaSuccess;
            if (ok) P->ev_in.push_back(a);
            ok = ok && cudaEventCreateWithFlags(&b, cudaEventDisableTiming) == cudaSuccess;
            if (ok) P->ev_k.push_back(b);
        }
        if (!ok) { wt_set_error("creating the pipeline's streams / events failed: %s", cudaGetErrorString(cudaGetLastError())); rc = WT_E_CUDA; }
    }
    if (rc) { wt_plan_destroy(P); return rc; }
    *plan_out = P;
    return WT_OK;
}

// H2D on s_in, kernels on the caller's stream, D2H on s_out; the caller's stream ends ordered after the last D2H
static int run_host_pipeline(WtPlan* P, cudaStream_t stream) {
    WT_CUDA(cudaEventRecord(P->ev_start, stream));
    WT_CUDA(cudaStreamWaitEvent(P->s_in, P->ev_start, 0));
    WT_CUDA(cudaStreamWaitEvent(P->s_out, P->ev_start, 0));
    const size_t G = P->children.size();
    for (size_t g = 0; g < G; ++g) {
        for (int i : P->child_layers[g])
            WT_CUDA(cudaMemcpyAsync(P->d_in[i], P->layers[i].in, (size_t)layer_numel(P->layers[i]) * sizeof(float),
                                    cudaMemcpyHostToDevice, P->s_in));
        WT_CUDA(cudaEventRecord(P->ev_in[g], P->s_in));
    }
    int launches = 0;
    for (size_t g = 0; g < G; ++g) {
        WT_CUDA(cudaStreamWaitEvent(stream, P->ev_in[g], 0));
        int rc = wt_plan_run(P->children[g], stream);
        if (rc) return rc;
        launches += P->children[g]->n_launches;
        WT_CUDA(cudaEventRecord(P->ev_k[g], stream));
        WT_CUDA(cudaStreamWaitEvent(P->s_out, P->ev_k[g], 0));
        for (int i : P->child_layers[g])
            WT_CUDA(cudaMemcpyAsync(P->layers[i].out, P->d_in[i], (size_t)layer_numel(P->layers[i]) * sizeof(float),
                                    cudaMemcpyDeviceToHost, P->s_out));
    }
    WT_CUDA(cudaEventRecord(P->ev_done, P->s_out));
    WT_CUDA(cudaStreamWaitEvent(stream, P->ev_done, 0));
    P->n_launches = launches;
    return WT_OK;
}

extern "C" int wt_plan_create(const WtLayerDesc* layers, int n_layers, int wavelet_id, int mode, int threshold_kind,
                              double threshold_value, int numpy_semantics, int host_io, WtPlan** plan_out) {
    WT_CHECK(plan_out != nullptr, WT_E_INVALID, "plan_out is NULL");
    *plan_out = nullptr;
    WT_CHECK(layers != nullptr || n_layers == 0, WT_E_INVALID, "layers is NULL");
    WT_CHECK(n_layers >= 0, WT_E_INVALID, "n_layers < 0");
    WT_CHECK(mode >= 0 && mode < WT_NUM_MODES, WT_E_INVALID, "Unknown mode name.");
    WT_CHECK(threshold_kind == WT_THRESHOLD_PERCENTILE || threshold_kind == WT_THRESHOLD_ABSOLUTE, WT_E_INVALID,
             "unknown threshold kind %d", threshold_kind);
    WT_CHECK(numpy_semantics == WT_NUMPY_LEGACY_1_22 || numpy_semantics == WT_NUMPY_NEP50_2X, WT_E_INVALID,
             "unknown numpy_semantics %d", numpy_semantics);
    if (threshold_kind == WT_THRESHOLD_PERCENTILE)
        WT_CHECK(threshold_value >= 0.0 && threshold_value <= 100.0, WT_E_INVALID,
                 "Percentiles must be in the range [0, 100]");
    if (host_io)
        return create_host_pipeline(layers, n_layers, wavelet_id, mode, threshold_kind, threshold_value, numpy_semantics,
                                    plan_out);
    WtPlan* P = new WtPlan();
    int rc = wt_get_bank(wavelet_id, &P->bank);
    if (rc) { delete P; return rc; }
    if (cudaGetDevice(&P->device) != cudaSuccess) {
        delete P;
        wt_set_error("no usable CUDA device (libwtprune has no CPU fallback): %s", cudaGetErrorString(cudaGetLastError()));
        return WT_E_CUDA;
    }
    P->wavelet_id = wavelet_id; P->mode = mode; P->thr_kind = threshold_kind; P->thr_value = threshold_value;
    P->sem = numpy_semantics; P->host_io = host_io;
    P->layers.assign(layers, layers + n_layers);
    P->path.assign(n_layers, PATH_EMPTY);
    P->d_in.assign(n_layers, nullptr); P->d_out.assign(n_layers, nullptr); P->d_coef.assign(n_layers, nullptr);
    P->ncoef.assign(n_layers, 0); P->zeros.assign(n_layers, 0);
    P->d_cand.assign(n_layers, nullptr); P->cand_cap.assign(n_layers, 0);
    const int F = P->bank.F;
    // Which layers take the fused recompute kernels (wt_fused.cuh) -- decided by measurement on the B200:
    //   absolute thresholds at level 0 (dwt_pruning_NoEntropy.py:35-45 on a <2-D tensor or with level 0): ONE pass over a
    //   whole model, 29 us against 118 us of per-layer launches for ResNet-18 (47 % of the measured HBM rate);
    //   small planes, absolute: the persistent CTAs hold few warps (80 registers, 68 KB of shared memory each) and lose to
    //   the one-tile-per-CTA kernels of wt_small.cu, which hide the HBM latency with occupancy (haar L1: 90 us vs 62 us);
    //   percentile thresholds: bit-exact and 1.1x the algorithmic DRAM traffic, but four latency-bound phases against the
    //   select's three on a 45 MB model and a slower streaming loop on large data (2.4 GB: 974 us vs 758 us).
    // So the round-1 kernels stay the default everywhere else.  WT_FUSED=all|abs|off overrides (all: every layer the fused
    // kernels support -- tests/test_gpu_fused.py keeps that path bit-exact; WT_LEGACY_PATH = off).
    int fused_mode = 1;
    if (const char* e = getenv("WT_FUSED")) fused_mode = !strcmp(e, "all") ? 2 : (!strcmp(e, "off") ? 0 : 1);
    if (getenv("WT_LEGACY_PATH")) fused_mode = 0;
    auto wants_fused = [&](const WtLayerDesc& L) {
        if (fused_mode == 2) return true;
        if (fused_mode == 0) return false;
        return threshold_kind == WT_THRESHOLD_ABSOLUTE && L.level == 0;
    };

    // ---- classify + size ----
    size_t bytes = 0;
    std::vector<size_t> off_coef(n_layers, 0), off_io(n_layers, 0), off_cand(n_layers, 0);
    long long tmp_elems = 0;
    int n_sel = 0;
    for (int i = 0; i < n_layers; ++i) {
        const WtLayerDesc& L = P->layers[i];
        if (L.batch < 0 || L.H < 0 || L.W < 0) { delete P; wt_set_error("layer %d: negative shape", i); return WT_E_INVALID; }
        if (L.level < 0) { delete P; wt_set_error("Level value of %d is too low . Minimum level is 0.", L.level); return WT_E_INVALID; }
        if (L.level > WT_MAX_LEVEL) { delete P; wt_set_error("layer %d: level %d above the supported maximum %d", i, L.level, WT_MAX_LEVEL); return WT_E_UNSUPPORTED; }
        const long long numel = layer_numel(L);
        if (numel == 0) continue;
        if ((L.in == nullptr || L.out == nullptr)) { delete P; wt_set_error("layer %d: NULL tensor pointer", i); return WT_E_INVALID; }
        if (wants_fused(L) && wt_fused_supported(L.H, L.W, F, L.level, mode) && numel * 4 < (1LL << 31)) {
            P->path[i] = PATH_FUSED;
            const int ch = L.level == 0 ? 0 : L.H, cw = L.level == 0 ? 0 : L.W;
            FusedClass* cls = nullptr;
            for (auto& c : P->fused) if (c.H == ch && c.W == cw) cls = &c;
            if (!cls) { FusedClass c; c.H = ch; c.W = cw; P->fused.push_back(c); cls = &P->fused.back(); }
            cls->layers.push_back(i);
        } else if (L.level == 0) {
            P->path[i] = PATH_LEVEL0;
        } else if (wt_small_supported(L.H, L.W, F, L.level, mode)) {
            P->path[i] = PATH_SMALL;
            bool found = false;
            for (auto& c : P->small) found |= (c.H == L.H && c.W == L.W);
            if (!found) { SmallClass c; c.H = L.H; c.W = L.W; P->small.push_back(c); }
        } else {
            P->path[i] = PATH_GENERIC;
            tmp_elems = std::max(tmp_elems, L.batch * wt_generic_tmp_elems(L.H, L.W, F, mode, L.level));
        }
        if (L.level != 0) {
            int sh, sw; int64_t nc;
            rc = wt_packed_shape(L.H, L.W, wavelet_id, L.level, mode, &sh, &sw, &nc);
            if (rc) { delete P; return rc; }
            P->ncoef[i] = L.batch * nc;
            P->zeros[i] = L.batch * ((long long)sh * sw - nc);
        }
        if (threshold_kind == WT_THRESHOLD_PERCENTILE && P->path[i] != PATH_FUSED) ++n_sel;
        if (P->path[i] != PATH_FUSED) P->legacy_zero = true;
    }
    // jobs of a fused class in ascending order of population (the ticket queue pipelines over them in this order)
    for (auto& c : P->fused)
        std::stable_sort(c.layers.begin(), c.layers.end(), [&](int a, int b) {
            const long long na = P->layers[a].level == 0 ? layer_numel(P->layers[a]) : P->ncoef[a];
            const long long nb = P->layers[b].level == 0 ? layer_numel(P->layers[b]) : P->ncoef[b];
            return na < nb;
        });
    {   // smallest class first: launch order of wt_plan_run
        auto work = [&](const FusedClass& c) { long long w = 0; for (int i : c.layers) w += layer_numel(P->layers[i]); return w; };
        std::stable_sort(P->fused.begin(), P->fused.end(), [&](const FusedClass& a, const FusedClass& b) { return work(a) < work(b); });
        long long all = 0;
        for (auto& c : P->fused) all += work(c);
        for (auto& c : P->fused) c.share = all ? (double)work(c) / (double)all : 1.0;
    }
    // ---- layout of the single device block ----
    auto take = [&](size_t sz) { size_t o = bytes; bytes = align_up(bytes + sz); return o; };
    const size_t off_sel_jobs = take((size_t)std::max(n_sel, 1) * sizeof(WtSelJob));
    std::vector<size_t> off_small(P->small.size());
    for (size_t c = 0; c < P->small.size(); ++c) off_small[c] = take((size_t)n_layers * sizeof(WtSmallJob));
    std::vector<size_t> off_fjobs(P->fused.size()), off_fscr(P->fused.size());
    for (size_t c = 0; c < P->fused.size(); ++c) {
        const size_t nj = P->fused[c].layers.size();
        off_fjobs[c] = take(nj * sizeof(WtFJob));
        P->fused[c].scratch_bytes = wt_fused_scratch_bytes((int)nj);
        off_fscr[c] = take(2 * P->fused[c].scratch_bytes);
    }
    const size_t off_thr = take((size_t)std::max(n_layers, 1) * sizeof(float));
    const size_t sel_scratch = wt_select_scratch_bytes(std::max(n_sel, 1));
    const size_t off_zero = take(sel_scratch + (size_t)std::max(n_layers, 1) * 4 * sizeof(long long));
    const size_t zero_bytes = bytes - off_zero;
    const size_t off_zero2 = take(zero_bytes);
    for (int i = 0; i < n_layers; ++i)
        if (P->path[i] == PATH_SMALL || P->path[i] == PATH_GENERIC) off_coef[i] = take((size_t)P->ncoef[i] * sizeof(float));
    if (threshold_kind == WT_THRESHOLD_PERCENTILE)
        for (int i = 0; i < n_layers; ++i)
            if (P->path[i] != PATH_EMPTY) {
                const long long pop = (P->layers[i].level == 0) ? layer_numel(P->layers[i]) : P->ncoef[i];
                P->cand_cap[i] = wt_select_cand_cap(pop);
                off_cand[i] = take((size_t)P->cand_cap[i] * sizeof(unsigned int));
            }
    int n_generic = 0;
    for (int i = 0; i < n_layers; ++i) n_generic += (P->path[i] == PATH_GENERIC);
    P->gen_lanes = std::max(1, std::min(n_generic, 4));
    std::vector<size_t> off_tmp0(P->gen_lanes), off_tmp1(P->gen_lanes);
    for (int q = 0; q < P->gen_lanes; ++q) {
        off_tmp0[q] = take((size_t)tmp_elems * sizeof(float));
        off_tmp1[q] = take((size_t)tmp_elems * sizeof(float));
    }
    if (host_io)
        for (int i = 0; i < n_layers; ++i)
            if (P->path[i] != PATH_EMPTY) off_io[i] = take((size_t)layer_numel(P->layers[i]) * sizeof(float));
    P->dev_bytes = bytes;
    cudaError_t e = cudaMalloc((void**)&P->dev, std::max(bytes, (size_t)256));
    if (e != cudaSuccess) {
        wt_set_error("cudaMalloc(%zu) failed: %s (libwtprune has no CPU fallback)", bytes, cudaGetErrorString(e));
        delete P;
        return WT_E_CUDA;
    }
    P->sel_jobs_dev = reinterpret_cast<WtSelJob*>(P->dev + off_sel_jobs);
    for (size_t c = 0; c < P->small.size(); ++c) P->small[c].jobs_dev = reinterpret_cast<WtSmallJob*>(P->dev + off_small[c]);
    for (size_t c = 0; c < P->fused.size(); ++c) {
        FusedClass& fc = P->fused[c];
        fc.jobs_dev = reinterpret_cast<WtFJob*>(P->dev + off_fjobs[c]);
        fc.scratch[0] = P->dev + off_fscr[c];
        fc.scratch[1] = fc.scratch[0] + fc.scratch_bytes;
        e = cudaMemset(fc.scratch[0], 0, 2 * fc.scratch_bytes);
        if (e != cudaSuccess) { wt_set_error("cudaMemset failed: %s", cudaGetErrorString(e)); wt_plan_destroy(P); return WT_E_CUDA; }
    }
    P->thresholds_dev = reinterpret_cast<float*>(P->dev + off_thr);
    P->zero_block = P->dev + off_zero;
    P->zero_bytes = zero_bytes;
    P->zero_blocks[0] = P->dev + off_zero;
    P->zero_blocks[1] = P->dev + off_zero2;
    P->counts_off = align_up(sel_scratch, 16);
    e = cudaMemset(P->zero_blocks[0], 0, zero_bytes);
    if (e == cudaSuccess) e = cudaMemset(P->zero_blocks[1], 0, zero_bytes);
    if (e != cudaSuccess) { wt_set_error("cudaMemset failed: %s", cudaGetErrorString(e)); wt_plan_destroy(P); return WT_E_CUDA; }
    wt_select_bind(&P->sel_args, P->zero_block, std::max(n_sel, 1));
    P->counts_dev = reinterpret_cast<long long*>(P->zero_block + align_up(sel_scratch, 16));
    P->sel_args.jobs = P->sel_jobs_dev;
    P->sel_args.n_jobs = n_sel;
    P->sel_args.thresholds = P->thresholds_dev;
    P->sel_args.counts = P->counts_dev;
    if (getenv("WT_SELECT_STAMPS")) {
        if (cudaMalloc((void**)&P->stamps_dev, (size_t)WT_SEL_STAMP_CTAS * 16 * 8) == cudaSuccess)
            cudaMemset(P->stamps_dev, 0, (size_t)WT_SEL_STAMP_CTAS * 16 * 8);
        P->sel_args.stamps = P->stamps_dev;
    }
    for (int q = 0; q < P->gen_lanes; ++q) {
        P->tmp0.push_back(reinterpret_cast<float*>(P->dev + off_tmp0[q]));
        P->tmp1.push_back(reinterpret_cast<float*>(P->dev + off_tmp1[q]));
    }
    for (int i = 0; i < n_layers; ++i) {
        if (P->path[i] == PATH_EMPTY) continue;
        if (P->path[i] == PATH_SMALL || P->path[i] == PATH_GENERIC) P->d_coef[i] = reinterpret_cast<float*>(P->dev + off_coef[i]);
        if (P->cand_cap[i]) P->d_cand[i] = reinterpret_cast<unsigned int*>(P->dev + off_cand[i]);
        if (host_io) {
            P->d_in[i] = P->d_out[i] = reinterpret_cast<float*>(P->dev + off_io[i]);
        } else {
            P->d_in[i] = const_cast<float*>(P->layers[i].in);
            P->d_out[i] = P->layers[i].out;
        }
    }
    const size_t n_streams = std::max(P->small.size() + P->fused.size(), (size_t)P->gen_lanes);
    if (n_streams > 1) {
        bool ok = cudaEventCreateWithFlags(&P->ev_fork, cudaEventDisableTiming) == cudaSuccess;
        for (size_t k = 1; ok && k < n_streams; ++k) {
            cudaStream_t q;
            cudaEvent_t ev;
            ok = cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking) == cudaSuccess;
            if (ok) P->aux.push_back(q);
            ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
            if (ok) P->ev_join.push_back(ev);
        }
        if (!ok) { wt_set_error("creating the plan's auxiliary streams failed: %s", cudaGetErrorString(cudaGetLastError())); wt_plan_destroy(P); return WT_E_CUDA; }
    }
    // WT_PRECOMPACT=1 (percentile plans with small-plane layers): sample pass + bracket + count / compact inside the analysis
    // kernels, so that the select never streams the spilled coefficients.  Bit-exact, and opt-in because it measured
    // slower than the select's own sampled pass at every model size tried: the ballots and the
    // staging double the analysis kernel and the two extra launches sit on the critical path.
    {
        const char* gr = getenv("WT_GRAPH");
        P->use_graph = (gr == nullptr || gr[0] == '\0') ? -1 : (gr[0] != '0' ? 1 : 0);
    }
    {
        const char* pre = getenv("WT_PRECOMPACT");
        P->precompact = threshold_kind == WT_THRESHOLD_PERCENTILE && !P->small.empty() && pre != nullptr && pre[0] != '\0' && pre[0] != '0';
    }
    rc = build_tables(P);
    if (!rc) rc = upload_tables(P);
    if (!rc) {
        // percentile plans: the select writes the thresholds; absolute plans: the constant, for wt_plan_thresholds_device
        std::vector<float> t0((size_t)std::max(n_layers, 1), threshold_kind == WT_THRESHOLD_ABSOLUTE ? (float)threshold_value : 0.0f);
        e = cudaMemcpy(P->thresholds_dev, t0.data(), t0.size() * sizeof(float), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { wt_set_error("cudaMemcpy failed: %s", cudaGetErrorString(e)); rc = WT_E_CUDA; }
    }
    if (rc) { wt_plan_destroy(P); return rc; }
    *plan_out = P;
    return WT_OK;
}

extern "C" size_t wt_plan_device_bytes(const WtPlan* P) {
    if (!P) return 0;
    size_t b = P->dev_bytes;
    for (const WtPlan* c : P->children) b += c->dev_bytes;
    return b;
}
extern "C" int wt_plan_num_launches(const WtPlan* P) { return P ? P->n_launches : 0; }

extern "C" int wt_plan_set_io(WtPlan* P, const WtLayerDesc* layers, int n_layers) {
    WT_CHECK(P != nullptr && layers != nullptr, WT_E_INVALID, "NULL argument");
    WT_CHECK(n_layers == (int)P->layers.size(), WT_E_INVALID, "wt_plan_set_io: layer count differs from the plan");
    for (int i = 0; i < n_layers; ++i) {
        const WtLayerDesc& a = P->layers[i];
        const WtLayerDesc& b = layers[i];
        WT_CHECK(a.batch == b.batch && a.H == b.H && a.W == b.W && a.level == b.level, WT_E_INVALID,
                 "wt_plan_set_io: layer %d shape/level differs from the plan", i);
        WT_CHECK(P->path[i] == PATH_EMPTY || (b.in && b.out), WT_E_INVALID, "layer %d: NULL tensor pointer", i);
    }
    for (int i = 0; i < n_layers; ++i) {
        P->layers[i].in = layers[i].in;
        P->layers[i].out = layers[i].out;
        if (!P->host_io && P->path[i] != PATH_EMPTY) {
            P->d_in[i] = const_cast<float*>(layers[i].in);
            P->d_out[i] = layers[i].out;
        }
    }
    if (P->host_io) return WT_OK;
    drop_graphs(P);
    int rc = build_tables(P);
    if (rc) return rc;
    return upload_tables(P);
}

extern "C" int wt_plan_set_peers(WtPlan* P, int n_peers, const int64_t* byte_deltas) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    WT_CHECK(n_peers >= 0 && n_peers <= WT_MAX_PEERS, WT_E_INVALID, "wt_plan_set_peers: 0..%d peers", WT_MAX_PEERS);
    WT_CHECK(n_peers == 0 || byte_deltas != nullptr, WT_E_INVALID, "wt_plan_set_peers: NULL deltas");
    WT_CHECK(!P->host_io, WT_E_UNSUPPORTED, "wt_plan_set_peers: host-I/O plans have no device output buffer to mirror");
    if (n_peers > 0) {
        // the kernels that mirror their stores: small-plane synthesis and the select's level-0 apply
        for (size_t i = 0; i < P->layers.size(); ++i) {
            WT_CHECK(P->path[i] != PATH_GENERIC, WT_E_UNSUPPORTED,
                     "wt_plan_set_peers: layer %d runs on the generic tiled path, which does not mirror its stores", (int)i);
            WT_CHECK(P->path[i] == PATH_FUSED || P->path[i] == PATH_EMPTY || P->thr_kind == WT_THRESHOLD_PERCENTILE, WT_E_UNSUPPORTED,
                     "wt_plan_set_peers: absolute-threshold plans mirror their stores on the fused path only");
        }
        for (int q = 0; q < n_peers; ++q)
            WT_CHECK((byte_deltas[q] & 15) == 0, WT_E_INVALID, "wt_plan_set_peers: peer mappings must keep 16-byte alignment");
    }
    P->peers.n = n_peers;
    P->peers.mc = 0;
    for (int q = 0; q < n_peers; ++q) P->peers.delta[q] = (long long)byte_deltas[q];
    drop_graphs(P);
    int rc = build_tables(P);
    if (rc) return rc;
    return upload_tables(P);
}

extern "C" int wt_plan_set_multicast(WtPlan* P, int64_t byte_delta) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    WT_CHECK(!P->host_io, WT_E_UNSUPPORTED, "wt_plan_set_multicast: host-I/O plans have no device output buffer to mirror");
    WT_CHECK(P->thr_kind == WT_THRESHOLD_PERCENTILE, WT_E_UNSUPPORTED, "wt_plan_set_multicast: percentile plans only");
    WT_CHECK((byte_delta & 15) == 0, WT_E_INVALID, "wt_plan_set_multicast: the multicast mapping must keep 16-byte alignment");
    for (size_t i = 0; i < P->layers.size(); ++i)
        WT_CHECK(P->path[i] != PATH_GENERIC && P->path[i] != PATH_FUSED, WT_E_UNSUPPORTED,
                 "wt_plan_set_multicast: layer %d runs on a path whose kernels do not store through multicast addresses", (int)i);
    memset(&P->peers, 0, sizeof(P->peers));
    P->peers.n = 1;
    P->peers.mc = 1;
    P->peers.delta[0] = (long long)byte_delta;
    drop_graphs(P);
    int rc = build_tables(P);
    if (rc) return rc;
    return upload_tables(P);
}

// the captured graphs hold kernel parameters: anything that rebuilds the tables or changes what a run launches drops them
static void drop_graphs(WtPlan* P) {
    for (int g = 0; g < 2; ++g) {
        if (P->graph_exec[g]) cudaGraphExecDestroy(P->graph_exec[g]);
        P->graph_exec[g] = nullptr;
    }
    P->graph_failed = 0;
}

// which of the two alternating scratch blocks the run being enqueued uses (host state only: a replayed graph needs it too)
static void advance_run_state(WtPlan* P) {
    // Percentile plans with a select launch enqueue no memset: two blocks alternate and the select kernel of this run
    // zeroes the block of the next one (every kernel that accumulates into the block runs in stream order after it).
    if (P->thr_kind == WT_THRESHOLD_PERCENTILE && !P->sel_jobs.empty()) {
        P->zero_block = P->zero_blocks[P->zpar];
        P->counts_dev = reinterpret_cast<long long*>(P->zero_block + P->counts_off);
        wt_select_rebind(&P->sel_args, P->zero_block, (int)P->sel_jobs.size());
        P->sel_args.counts = P->counts_dev;
        P->sel_args.zero_next = P->zero_blocks[P->zpar ^ 1];
        P->sel_args.zero_next_bytes = P->zero_bytes;
        P->zpar ^= 1;
    }
    for (auto& c : P->fused) {
        c.run_parity = c.parity;
        c.parity ^= 1;
    }
}

static int plan_enqueue(WtPlan* P, cudaStream_t stream) {
    const int n = (int)P->layers.size();
    int launches = 0;
    P->n_phases = 0;
    int mark_rc = WT_OK, mark_launches = 0;
    // closes the previous phase (launch count) and opens phase `name` (nullptr = just close)
    auto mark = [&](const char* name) {
        if (!P->timing || mark_rc) return;
        if (P->n_phases) P->phase_launches[P->n_phases - 1] = launches - mark_launches;
        mark_launches = launches;
        const size_t slot = P->n_phases;
        if (P->events.size() <= slot) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) { mark_rc = WT_E_CUDA; return; }
            P->events.push_back(e);
            P->phase_names.push_back("");
            P->phase_launches.push_back(0);
        }
        if (cudaEventRecord(P->events[slot], stream) != cudaSuccess) { mark_rc = WT_E_CUDA; return; }
        if (name) { P->phase_names[slot] = name; ++P->n_phases; }
    };
    char label[64];
    // counts / select scratch of the launches that accumulate with atomics (generic path, legacy kernels); the fused
    // classes zero their own scratch for the next run and store their counts, so a fused-only plan enqueues no memset
    // (percentile plans with a select launch: advance_run_state)
    const bool self_zero = P->thr_kind == WT_THRESHOLD_PERCENTILE && !P->sel_jobs.empty();
    const size_t counts_off = P->counts_off;
    const bool split_zero = !self_zero && P->legacy_zero && !P->timing && !P->aux.empty() && counts_off > 0 && counts_off < P->zero_bytes;
    if (P->legacy_zero && !self_zero) {
        mark("memset");
        if (split_zero) WT_CUDA(cudaMemsetAsync(P->zero_block + counts_off, 0, P->zero_bytes - counts_off, stream));
        else WT_CUDA(cudaMemsetAsync(P->zero_block, 0, P->zero_bytes, stream));
    }
    int rc;
    // 1. analysis
    // size classes fork onto the auxiliary streams (kept serial on the caller's stream while phases are timed)
    const bool forked = !P->timing && !P->aux.empty();
    auto fork = [&]() -> int {
        if (!forked) return WT_OK;
        WT_CUDA(cudaEventRecord(P->ev_fork, stream));
        for (cudaStream_t q : P->aux) WT_CUDA(cudaStreamWaitEvent(q, P->ev_fork, 0));
        return WT_OK;
    };
    auto join = [&]() -> int {
        if (!forked) return WT_OK;
        for (size_t k = 0; k < P->aux.size(); ++k) {
            WT_CUDA(cudaEventRecord(P->ev_join[k], P->aux[k]));
            WT_CUDA(cudaStreamWaitEvent(stream, P->ev_join[k], 0));
        }
        return WT_OK;
    };
    auto class_stream = [&](size_t k) { return (forked && k > 0) ? P->aux[k - 1] : stream; };
    if ((rc = fork())) return rc;
    if (split_zero) WT_CUDA(cudaMemsetAsync(P->zero_block, 0, counts_off, P->aux.back()));
    // 0. the fused classes: one cooperative launch each, independent of everything below.  Smallest class first, each
    // on its own stream; a later (larger) class leaves room for the CTAs of the earlier ones, so that the kernels are
    // resident together (if the guess is wrong the larger launch simply waits for the smaller kernels to end).
    int reserve = 0;
    for (size_t k = 0; k < P->fused.size(); ++k) {
        FusedClass& c = P->fused[k];
        if (c.total_items <= 0) continue;
        if (c.H) snprintf(label, sizeof(label), "fused_%dx%d", c.H, c.W); else snprintf(label, sizeof(label), "fused_flat");
        mark(label);
        WtFArgs fa;
        memset(&fa, 0, sizeof(fa));
        fa.jobs = c.jobs_dev;
        fa.n_jobs = (int)c.jobs.size(); fa.n_select = c.n_select;
        fa.total_items = c.total_items; fa.total_s_items = c.total_s_items;
        fa.scratch = c.scratch[c.run_parity]; fa.scratch_next = c.scratch[c.run_parity ^ 1]; fa.scratch_bytes = c.scratch_bytes;
        fa.thresholds = P->thresholds_dev; fa.counts = P->counts_dev;
        cudaStream_t cs = class_stream(P->small.size() + k);
        if (getenv("WT_FUSED_STAMPS")) {
            const size_t sb = (size_t)WT_SEL_STAMP_CTAS * WT_F_STAMP_SLOTS * 8;
            if (!c.stamps_dev && cudaMalloc((void**)&c.stamps_dev, sb) != cudaSuccess) c.stamps_dev = nullptr;
            if (c.stamps_dev) cudaMemsetAsync(c.stamps_dev, 0, sb, cs);
            fa.stamps = c.stamps_dev;
        }
        int grid = 0;
        rc = wt_fused_launch(c.H, c.W, P->bank, P->arithmetic, fa, reserve, k + 1 == P->fused.size() ? 1.0 : c.share, cs, &grid);
        if (rc) return rc;
        reserve += 2 * grid;
        ++launches;
    }
    const bool pre = P->precompact && !P->sel_jobs.empty();
    if (pre) {
        // the select's S0 on the analysis side: 12-bit histogram of one tile in 2^shift, per class; then ONE small kernel
        // turns every job's sample histogram into its bracket (all classes' histograms are complete: join first)
        for (size_t k = 0; k < P->small.size(); ++k) {
            auto& c = P->small[k];
            snprintf(label, sizeof(label), "small_sample_%dx%d", c.H, c.W);
            mark(label);
            rc = wt_small_sample_launch(c.H, c.W, P->bank, P->arithmetic, c.jobs_dev, (int)c.jobs.size(), c.n_tiles, P->sel_args.hist1,
                                        P->sel_args.hist1c, class_stream(k));
            if (rc) return rc;
            if (c.n_tiles) ++launches;
        }
        if ((rc = join())) return rc;
        mark("select_bracket");
        rc = wt_select_bracket_launch(P->sel_args, stream);
        if (rc) return rc;
        ++launches;
        if ((rc = fork())) return rc;
    }
    for (size_t k = 0; k < P->small.size(); ++k) {
        auto& c = P->small[k];
        snprintf(label, sizeof(label), "small_analysis_%dx%d", c.H, c.W);
        mark(label);
        if (pre)
            rc = wt_small_analysis_compact_launch(c.H, c.W, P->bank, P->arithmetic, c.jobs_dev, (int)c.jobs.size(), c.n_tiles,
                                                  P->counts_dev, P->sel_args.state, class_stream(k));
        else
            rc = wt_small_analysis_launch(c.H, c.W, P->bank, P->arithmetic, c.jobs_dev, (int)c.jobs.size(), c.n_tiles, P->counts_dev,
                                          class_stream(k));
        if (rc) return rc;
        if (c.n_tiles) ++launches;
    }
    int gi = 0;
    for (auto& g : P->generic) {
        snprintf(label, sizeof(label), "generic_analysis_layer%d", g.layer);
        mark(label);
        rc = (P->arithmetic ? wt_generic_analysis_fma : wt_generic_analysis)(P->bank, P->mode, g, P->counts_dev,
                                                                             class_stream((size_t)(gi++ % P->gen_lanes)), &launches);
        if (rc) return rc;
    }
    if ((rc = join())) return rc;
    // 2. threshold: exact percentile select (+ fused level-0 apply), or plain compare
    if (P->thr_kind == WT_THRESHOLD_PERCENTILE) {
        if (!P->sel_jobs.empty()) {
            mark(P->sel_args.any_apply ? "select_apply" : "select");
            rc = wt_select_launch(P->sel_args, stream);
            if (rc) return rc;
            ++launches;
        }
    } else {
        if (!P->level0_abs.empty()) mark("threshold_level0");
        for (int i : P->level0_abs) {
            rc = wt_apply_threshold_launch(P->d_in[i], P->d_out[i], layer_numel(P->layers[i]), nullptr, (float)P->thr_value,
                                           P->counts_dev + (size_t)i * 4, stream);
            if (rc) return rc;
            ++launches;
        }
    }
    // 3. synthesis with the threshold applied on load
    if ((rc = fork())) return rc;
    for (size_t k = 0; k < P->small.size(); ++k) {
        auto& c = P->small[k];
        snprintf(label, sizeof(label), "small_synthesis_%dx%d", c.H, c.W);
        mark(label);
        rc = wt_small_synthesis_launch(c.H, c.W, P->bank, P->arithmetic, c.jobs_dev, (int)c.jobs.size(), c.n_tiles, P->thresholds_dev,
                                       P->counts_dev, class_stream(k));
        if (rc) return rc;
        if (c.n_tiles) ++launches;
    }
    gi = 0;
    for (auto& g : P->generic) {
        snprintf(label, sizeof(label), "generic_synthesis_layer%d", g.layer);
        mark(label);
        rc = (P->arithmetic ? wt_generic_synthesis_fma : wt_generic_synthesis)(P->bank, P->mode, g, P->thresholds_dev, P->counts_dev, class_stream((size_t)(gi++ % P->gen_lanes)),
                                  &launches);
        if (rc) return rc;
    }
    if ((rc = join())) return rc;
    if (P->host_io) mark("d2h");
    if (P->host_io)
        for (int i = 0; i < n; ++i)
            if (P->path[i] != PATH_EMPTY)
                WT_CUDA(cudaMemcpyAsync(P->layers[i].out, P->d_out[i], (size_t)layer_numel(P->layers[i]) * sizeof(float),
                                        cudaMemcpyDeviceToHost, stream));
    mark(nullptr);
    WT_CHECK(mark_rc == WT_OK, WT_E_CUDA, "wt_plan_run: cudaEvent failure while timing phases");
    P->n_launches = launches;
    return WT_OK;
}

extern "C" int wt_plan_run(WtPlan* P, void* stream_) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (P->host_io) return run_host_pipeline(P, stream);
    const int par = (int)(P->run_count & 1ull);   // both alternations flip once per run
    advance_run_state(P);
    ++P->run_count;
    // automatic: from its second run on, a plan of several dependent launches (with the forks onto the class streams and the
    // joins in between) replays them as a graph: 1.5-2 % on the conv-weight plans (7 launches), 2-7 % on large-plane plans
    // (a dozen short launches per plane); single-launch plans (the level-0 flows) gain nothing and stay on plain launches
    const bool want_graph = P->use_graph > 0 || (P->use_graph < 0 && P->n_launches >= WT_GRAPH_AUTO_MIN_LAUNCHES);
    const bool graphable = want_graph && !P->graph_failed && !P->timing && P->stamps_dev == nullptr && getenv("WT_FUSED_STAMPS") == nullptr;
    if (!graphable) return plan_enqueue(P, stream);
    if (!P->graph_exec[par]) {
        // capture this parity's launch sequence (the class streams fork from and join back into `stream`, so they are
        // branches of the same graph); nothing runs during the capture, the replay below is the run
        if (cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();   // e.g. the legacy default stream cannot be captured
            P->graph_failed = 1;
            return plan_enqueue(P, stream);
        }
        const int rc = plan_enqueue(P, stream);
        cudaGraph_t g = nullptr;
        const cudaError_t e = cudaStreamEndCapture(stream, &g);
        if (rc || e != cudaSuccess || g == nullptr) {
            if (g) cudaGraphDestroy(g);
            cudaGetLastError();
            P->graph_failed = 1;
            if (rc) return rc;
            return plan_enqueue(P, stream);
        }
        cudaGraphExec_t x = nullptr;
        const cudaError_t ei = cudaGraphInstantiate(&x, g, 0);
        cudaGraphDestroy(g);
        if (ei != cudaSuccess || x == nullptr) {
            cudaGetLastError();
            P->graph_failed = 1;
            return plan_enqueue(P, stream);
        }
        P->graph_exec[par] = x;
        P->graph_launches[par] = P->n_launches;
    }
    WT_CUDA(cudaGraphLaunch(P->graph_exec[par], stream));
    P->n_launches = P->graph_launches[par];
    return WT_OK;
}

// ---- staged runs: analysis and synthesis as separate calls, the threshold supplied by the caller ----------------------
// For callers that take the order statistic OUTSIDE the plan -- the intra-layer multi-GPU split (sharded.py:
// prune_tensor_split) agrees on one threshold over every rank's coefficients.  Stage 1 transforms (coefficients stay in
// the plan's workspace: wt_plan_layer_coefficients), stage 2 thresholds with whatever sits in the plan's threshold
// array (wt_plan_thresholds_device) and reconstructs.  Percentile plans on the caller's stream, kernels one after the other.
extern "C" int wt_plan_run_stage(WtPlan* P, int stage, void* stream_) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    WT_CHECK(stage == 1 || stage == 2, WT_E_INVALID, "wt_plan_run_stage: stage 1 (analysis) or 2 (threshold + synthesis)");
    WT_CHECK(!P->host_io && P->thr_kind == WT_THRESHOLD_PERCENTILE && P->fused.empty() && !P->precompact && P->peers.n == 0,
             WT_E_UNSUPPORTED, "wt_plan_run_stage: device percentile plans on the spill kernels, without peers");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int n = (int)P->layers.size();
    int launches = 0, rc = WT_OK;
    if (stage == 1) {
        advance_run_state(P);
        ++P->run_count;
        // no select kernel runs here, and it is the select that zeroes the block of the plan's next run: clear both, so that
        // staged and whole runs of one plan can be mixed
        for (int q = 0; q < 2; ++q)
            if (P->zero_blocks[q] && P->zero_bytes) WT_CUDA(cudaMemsetAsync(P->zero_blocks[q], 0, P->zero_bytes, stream));
        for (auto& c : P->small) {
            rc = wt_small_analysis_launch(c.H, c.W, P->bank, P->arithmetic, c.jobs_dev, (int)c.jobs.size(), c.n_tiles, P->counts_dev, stream);
            if (rc) return rc;
            if (c.n_tiles) ++launches;
        }
        for (auto& g : P->generic) {
            rc = (P->arithmetic ? wt_generic_analysis_fma : wt_generic_analysis)(P->bank, P->mode, g, P->counts_dev, stream, &launches);
            if (rc) return rc;
        }
    } else {
        for (int i = 0; i < n; ++i)   // level-0 layers: the compare itself (it also counts the nonzero inputs)
            if (P->path[i] == PATH_LEVEL0) {
                rc = wt_apply_threshold_launch(P->d_in[i], P->d_out[i], layer_numel(P->layers[i]), P->thresholds_dev + i, 0.0f,
                                               P->counts_dev + (size_t)i * 4, stream);
                if (rc) return rc;
                ++launches;
            }
        for (auto& c : P->small) {
            rc = wt_small_synthesis_launch(c.H, c.W, P->bank, P->arithmetic, c.jobs_dev, (int)c.jobs.size(), c.n_tiles, P->thresholds_dev,
                                           P->counts_dev, stream);
            if (rc) return rc;
            if (c.n_tiles) ++launches;
        }
        for (auto& g : P->generic) {
            rc = (P->arithmetic ? wt_generic_synthesis_fma : wt_generic_synthesis)(P->bank, P->mode, g, P->thresholds_dev, P->counts_dev, stream,
                                                                                   &launches);
            if (rc) return rc;
        }
    }
    P->n_launches = launches;
    return WT_OK;
}

extern "C" int wt_plan_layer_coefficients(WtPlan* P, int layer, const float** coef, int64_t* n_explicit, int64_t* n_implicit_zeros) {
    WT_CHECK(P != nullptr && coef != nullptr && n_explicit != nullptr && n_implicit_zeros != nullptr, WT_E_INVALID, "NULL argument");
    WT_CHECK(!P->host_io, WT_E_UNSUPPORTED, "host-I/O plans keep their workspaces per layer group");
    WT_CHECK(layer >= 0 && layer < (int)P->layers.size(), WT_E_INVALID, "layer %d out of range", layer);
    *coef = nullptr; *n_explicit = 0; *n_implicit_zeros = 0;
    switch (P->path[layer]) {
        case PATH_EMPTY: return WT_OK;
        case PATH_LEVEL0: *coef = P->d_in[layer]; *n_explicit = layer_numel(P->layers[layer]); return WT_OK;   // level 0: the weights themselves
        case PATH_SMALL:
        case PATH_GENERIC: *coef = P->d_coef[layer]; *n_explicit = P->ncoef[layer]; *n_implicit_zeros = P->zeros[layer]; return WT_OK;
        default: wt_set_error("layer %d runs on the fused recompute kernels, which keep no coefficients", layer); return WT_E_UNSUPPORTED;
    }
}

extern "C" int wt_plan_set_graph(WtPlan* P, int enable) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    for (WtPlan* c : P->children) {
        int rc = wt_plan_set_graph(c, enable);
        if (rc) return rc;
    }
    P->use_graph = enable ? 1 : 0;
    if (!enable) drop_graphs(P);
    return WT_OK;
}

extern "C" int wt_plan_set_arithmetic(WtPlan* P, int arithmetic) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    WT_CHECK(arithmetic == WT_ARITH_FAITHFUL || arithmetic == WT_ARITH_FMA_TAPORDER, WT_E_INVALID, "unknown arithmetic %d", arithmetic);
    if (arithmetic == WT_ARITH_FMA_TAPORDER) {
#ifndef WT_WITH_FMA_KERNELS
        for (const auto& c : P->fused)
            WT_CHECK(c.H == 0, WT_E_UNSUPPORTED, "fma_taporder: this build has no fused small-plane kernels in that arithmetic "
                                                 "(WT_BUILD_FMA=1); the default routing does not use them");
#endif
    }
    for (WtPlan* c : P->children) {
        int rc = wt_plan_set_arithmetic(c, arithmetic);
        if (rc) return rc;
    }
    P->arithmetic = arithmetic;
    drop_graphs(P);
    return WT_OK;
}

extern "C" int wt_plan_select_path(const WtPlan* P) {
    if (!P) return -1;
    if (P->host_io) return P->children.empty() ? -1 : wt_plan_select_path(P->children[0]);
    if (P->thr_kind != WT_THRESHOLD_PERCENTILE) return -1;
    for (const auto& c : P->fused) if (c.n_select > 0) return 2;
    if (P->sel_jobs.empty()) return -1;
    if (P->precompact) return 3;
    return wt_select_path(P->sel_args.n_tiles, P->sel_args.short_lists);
}

extern "C" int wt_plan_set_timing(WtPlan* P, int enable) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    P->timing = enable ? 1 : 0;
    if (!enable) P->n_phases = 0;
    return WT_OK;
}

extern "C" int wt_plan_phase_times(WtPlan* P, int max_phases, float* ms, int* launches, char* names, int name_stride) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    const int n = (int)P->n_phases;
    if (n == 0) return 0;
    WT_CUDA(cudaEventSynchronize(P->events[n]));
    for (int i = 0; i < n && i < max_phases; ++i) {
        float t = 0.f;
        WT_CUDA(cudaEventElapsedTime(&t, P->events[i], P->events[i + 1]));
        if (ms) ms[i] = t;
        if (launches) launches[i] = P->phase_launches[i];
        if (names && name_stride > 0) {
            strncpy(names + (size_t)i * name_stride, P->phase_names[i].c_str(), name_stride - 1);
            names[(size_t)i * name_stride + name_stride - 1] = 0;
        }
    }
    return n < max_phases ? n : max_phases;
}

extern "C" int wt_plan_counts_device(WtPlan* P, const int64_t** counts_dev) {
    WT_CHECK(P != nullptr && counts_dev != nullptr, WT_E_INVALID, "NULL argument");
    WT_CHECK(!P->host_io, WT_E_UNSUPPORTED, "host-I/O plans keep their counts per layer group: use wt_plan_results");
    *counts_dev = reinterpret_cast<const int64_t*>(P->counts_dev);
    return WT_OK;
}

extern "C" int wt_plan_thresholds_device(WtPlan* P, const float** thresholds_dev) {
    WT_CHECK(P != nullptr && thresholds_dev != nullptr, WT_E_INVALID, "NULL argument");
    WT_CHECK(!P->host_io, WT_E_UNSUPPORTED, "host-I/O plans keep their thresholds per layer group: use wt_plan_results");
    *thresholds_dev = P->thresholds_dev;
    return WT_OK;
}

extern "C" int wt_plan_results(WtPlan* P, void* stream_, int64_t* counts, int64_t* nonzero_before, float* thresholds) {
    WT_CHECK(P != nullptr, WT_E_INVALID, "plan is NULL");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int n = (int)P->layers.size();
    if (P->host_io) {
        for (int i = 0; i < n; ++i) {
            if (counts) { counts[i * 3 + 0] = layer_numel(P->layers[i]); counts[i * 3 + 1] = counts[i * 3 + 2] = 0; }
            if (nonzero_before) nonzero_before[i] = 0;
            if (thresholds) thresholds[i] = (P->thr_kind == WT_THRESHOLD_ABSOLUTE) ? (float)P->thr_value : 0.0f;
        }
        for (size_t g = 0; g < P->children.size(); ++g) {
            const std::vector<int>& idx = P->child_layers[g];
            std::vector<int64_t> c(idx.size() * 3), nb(idx.size());
            std::vector<float> th(idx.size());
            int rc = wt_plan_results(P->children[g], stream_, c.data(), nb.data(), th.data());
            if (rc) return rc;
            for (size_t k = 0; k < idx.size(); ++k) {
                if (counts) for (int q = 0; q < 3; ++q) counts[idx[k] * 3 + q] = c[k * 3 + q];
                if (nonzero_before) nonzero_before[idx[k]] = nb[k];
                if (thresholds) thresholds[idx[k]] = th[k];
            }
        }
        return WT_OK;
    }
    std::vector<long long> c4((size_t)n * 4);
    std::vector<float> thr(n);
    if (n) {
        WT_CUDA(cudaMemcpyAsync(c4.data(), P->counts_dev, c4.size() * sizeof(long long), cudaMemcpyDeviceToHost, stream));
        WT_CUDA(cudaMemcpyAsync(thr.data(), P->thresholds_dev, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, stream));
    }
    WT_CUDA(cudaStreamSynchronize(stream));
    if (getenv("WT_FUSED_DEBUG")) {   // tuning aid: the last run of every fused class (failed jobs, phase boundaries per CTA)
        for (auto& c : P->fused) {
            const char* scr = c.scratch[c.parity ^ 1];
            WtFHeader h;
            WT_CUDA(cudaMemcpy(&h, scr, sizeof(h), cudaMemcpyDeviceToHost));
            fprintf(stderr, "[wt fused %dx%d] jobs %zu items %d sample items %d failed %u wide %u\n", c.H, c.W, c.jobs.size(), c.total_items,
                    c.total_s_items, h.fail_count, h.wide);
            if (c.stamps_dev) {
                const int SL = WT_F_STAMP_SLOTS;
                std::vector<unsigned long long> sp((size_t)WT_SEL_STAMP_CTAS * SL);
                WT_CUDA(cudaMemcpy(sp.data(), c.stamps_dev, sp.size() * 8, cudaMemcpyDeviceToHost));
                unsigned long long t0 = ~0ull;
                int g = 0;
                for (int q = 0; q < WT_SEL_STAMP_CTAS; ++q) if (sp[(size_t)q * SL]) { t0 = std::min(t0, sp[(size_t)q * SL]); g = q + 1; }
                double mx[8] = {0}, sum[8] = {0};
                for (int q = 0; q < g; ++q)
                    for (int k = 0; k < 7; ++k) {
                        const unsigned long long raw = sp[(size_t)q * SL + k];
                        const double v = raw ? (double)(raw - t0) * 1e-3 : 0.0;
                        mx[k] = std::max(mx[k], v);
                        sum[k] += v / g;
                    }
                fprintf(stderr, "  %d CTAs, us mean/max: start %.1f/%.1f endS %.1f/%.1f endC %.1f/%.1f endT1 %.1f/%.1f startA %.1f/%.1f endA %.1f/%.1f endX %.1f/%.1f\n",
                        g, sum[0], mx[0], sum[1], mx[1], sum[2], mx[2], sum[3], mx[3], sum[4], mx[4], sum[5], mx[5], sum[6], mx[6]);
            }
            std::vector<WtFState> st(c.jobs.size());
            WT_CUDA(cudaMemcpy(st.data(), scr + sizeof(WtFHeader), st.size() * sizeof(WtFState), cudaMemcpyDeviceToHost));
            for (size_t q = 0; q < st.size(); ++q)
                if (st[q].failed || getenv("WT_FUSED_DEBUG_ALL"))
                    fprintf(stderr, "  job %zu layer %d n %lld zeros %lld k0 %lld shift %d items %d/%d: lo %u width %u below %llu cand %u (cap %u) a %g b %g t %g%s\n",
                            q, c.jobs[q].layer, c.jobs[q].n, c.jobs[q].zeros, c.jobs[q].k0, c.jobs[q].sample_shift, c.jobs[q].items,
                            c.jobs[q].s_items, st[q].lo_key, st[q].width, st[q].below, st[q].s_count, c.jobs[q].cand_cap, st[q].a, st[q].b,
                            st[q].t, st[q].failed ? "  FAILED -> exact path" : "");
        }
    }
    if (P->stamps_dev && !P->sel_jobs.empty()) {
        std::vector<unsigned long long> st((size_t)WT_SEL_STAMP_CTAS * 16);
        WT_CUDA(cudaMemcpy(st.data(), P->stamps_dev, st.size() * 8, cudaMemcpyDeviceToHost));
        WT_CUDA(cudaMemset(P->stamps_dev, 0, st.size() * 8));
        unsigned long long t0 = ~0ull;
        int g = 0;
        for (int c = 0; c < WT_SEL_STAMP_CTAS; ++c) if (st[c * 16]) { t0 = std::min(t0, st[c * 16]); g = c + 1; }
        double mx[16] = {0}, sum[16] = {0};
        for (int c = 0; c < g; ++c)
            for (int k = 0; k < 16; ++k) {
                double v = (k < 7) ? (double)(st[c * 16 + k] - t0) * 1e-3 : (double)st[c * 16 + k] * 1e-3;
                if (k < 7 && st[c * 16 + k] == 0) v = 0;  // phase not run (radix-only path)
                mx[k] = std::max(mx[k], v);
                sum[k] += v / g;
            }
        fprintf(stderr, "[wt select %d CTAs, us, mean/max] endS0 %.1f/%.1f endS1 %.1f/%.1f endC1 %.1f/%.1f endC2 %.1f/%.1f startP4 %.1f/%.1f end %.1f/%.1f"
                " || acc: sample %.1f/%.1f flushS %.1f/%.1f scanS %.1f/%.1f stream %.1f/%.1f C1 %.1f/%.1f C2 %.1f/%.1f scan4 %.1f/%.1f apply %.1f/%.1f radix %.1f/%.1f\n",
                g, sum[1], mx[1], sum[2], mx[2], sum[3], mx[3], sum[4], mx[4], sum[5], mx[5], sum[6], mx[6], sum[7], mx[7],
                sum[8], mx[8], sum[9], mx[9], sum[10], mx[10], sum[11], mx[11], sum[12], mx[12], sum[13], mx[13], sum[14], mx[14],
                sum[15], mx[15]);
        if (getenv("WT_SELECT_STAMPS_CTAS")) {  // per-CTA arrival at the barriers (us): who is everybody waiting for
            for (int c = 0; c < g; ++c) {
                fprintf(stderr, "  cta %3d:", c);
                for (int k = 0; k < 7; ++k) fprintf(stderr, " %6.1f", st[c * 16 + k] ? (double)(st[c * 16 + k] - t0) * 1e-3 : 0.0);
                fprintf(stderr, "\n");
            }
        }
        std::vector<WtSelState> ss(P->sel_jobs.size());
        WT_CUDA(cudaMemcpy(ss.data(), P->sel_args.state, ss.size() * sizeof(WtSelState), cudaMemcpyDeviceToHost));
        for (size_t q = 0; q < ss.size(); ++q) {
            const WtSelJob& jb = P->sel_jobs[q];
            if (ss[q].failed || getenv("WT_SELECT_STAMPS_ALL"))
                fprintf(stderr, "  job %zu layer %d n %lld zeros %lld k0 %lld shift %d: bracket %u+%u bins, below %llu, in bracket %u (cap %u), a %g b %g%s",
                        q, jb.layer, jb.n, jb.zeros, jb.k0, jb.sample_shift, ss[q].lo_bin, ss[q].n_bins, ss[q].below,
                        ss[q].s_count, jb.cand_cap, ss[q].a, ss[q].b, ss[q].failed ? "  -> RADIX PATH" : "");
            if (ss[q].failed && ss[q].n_bins == 0) fprintf(stderr, " (sample bins %u..%u)", ss[q].cprefix[0], ss[q].cprefix[1]);
            if (ss[q].failed || getenv("WT_SELECT_STAMPS_ALL")) fprintf(stderr, "\n");
        }
    }
    for (int i = 0; i < n; ++i) {
        if (counts) {
            counts[i * 3 + 0] = layer_numel(P->layers[i]);
            counts[i * 3 + 1] = c4[(size_t)i * 4 + 1];
            counts[i * 3 + 2] = c4[(size_t)i * 4 + 2];
        }
        if (nonzero_before) nonzero_before[i] = c4[(size_t)i * 4 + 3];
        if (thresholds) thresholds[i] = (P->thr_kind == WT_THRESHOLD_ABSOLUTE) ? (float)P->thr_value : thr[i];
    }
    return WT_OK;
}

extern "C" int wt_dwt_prune_f32(const WtLayerDesc* layers, int n_layers, int wavelet_id, int mode, int threshold_kind,
                                double threshold_value, int numpy_semantics, void* stream, int64_t* counts,
                                int64_t* nonzero_before, float* thresholds) {
    WtPlan* P = nullptr;
    int rc = wt_plan_create(layers, n_layers, wavelet_id, mode, threshold_kind, threshold_value, numpy_semantics, 0, &P);
    if (rc) return rc;
    rc = wt_plan_run(P, stream);
    if (!rc) rc = wt_plan_results(P, stream, counts, nonzero_before, thresholds);
    else cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
    wt_plan_destroy(P);
    return rc;
}
