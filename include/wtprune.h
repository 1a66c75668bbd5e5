/*
 * wtprune.h -- C ABI of libwtprune.so, the B200 (sm_100a) implementation of the reference's
 * DWT-prune hot path and its two sibling maskers.
 *
 * Every entry point replaces one interface of iAmGiG/WaveletTransforms (paths relative to the
 * reference checkout) or of the third-party code that interface calls (PyWavelets / NumPy / ATen,
 * which are not part of the reference tree).  Plain pointers and sizes only: no torch types.
 *
 * Conventions
 *   - every function returns 0 on success, a negative WT_E_* code on failure; wt_last_error()
 *     returns a thread-local message for the last failure on the calling thread;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream);
 *   - device pointers unless a parameter is documented as HOST;
 *   - all tensors are contiguous row-major fp32; a "plane" is the last two axes (H x W) and
 *     `batch` is the product of all leading axes (conv weight [O,I,kH,kW] -> batch O*I, H kH, W kW);
 *   - the library never falls back to the CPU: without a usable CUDA device every compute entry
 *     point fails with WT_E_CUDA.
 */
#ifndef WTPRUNE_H
#define WTPRUNE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WT_OK 0
#define WT_E_INVALID (-1)   /* bad argument (unknown wavelet / mode, level < 0, null pointer ...) */
#define WT_E_CUDA (-2)      /* CUDA runtime failure (message in wt_last_error) */
#define WT_E_WORKSPACE (-3) /* caller workspace too small */
#define WT_E_UNSUPPORTED (-4)

/* pywt boundary modes on the reference's path: dwt_pruning.py:35 (default 'periodization');
 * 'symmetric' is pywt's own default and the second mode carried. */
#define WT_MODE_PERIODIZATION 0
#define WT_MODE_SYMMETRIC 1
/* the rest of pywt.Modes.modes -- `mode` is an open string parameter of multi_resolution_analysis
 * (dwt_pruning.py:35, dwt_pruning_NoEntropy.py:12); they share the symmetric mode's coefficient
 * lengths and synthesis and differ in the analysis-side signal extension only */
#define WT_MODE_ZERO 2
#define WT_MODE_CONSTANT 3
#define WT_MODE_REFLECT 4
#define WT_MODE_PERIODIC 5
#define WT_MODE_SMOOTH 6
#define WT_MODE_ANTISYMMETRIC 7
#define WT_MODE_ANTIREFLECT 8
#define WT_NUM_MODES 9

/* np.percentile arithmetic: the reference pins numpy 1.22.4 (requirements.txt:2),
 * the GPU box runs numpy 2.x. */
#define WT_NUMPY_LEGACY_1_22 0
#define WT_NUMPY_NEP50_2X 1

/* threshold kinds */
#define WT_THRESHOLD_PERCENTILE 0 /* dwt_pruning.py:25-32 */
#define WT_THRESHOLD_ABSOLUTE 1   /* dwt_pruning_NoEntropy.py:45 */

typedef struct WtLayerDesc {
    const float* in; /* weight tensor [batch][H][W] (device; HOST for host-I/O plans) */
    float* out;      /* pruned tensor, same shape; may alias `in` */
    int64_t batch;   /* product of leading axes; a <2-D tensor is batch=1,H=1,W=numel with level 0 */
    int32_t H, W;
    int32_t level;   /* decomposition level actually applied to this tensor (>= 0); 0 = the
                        select-and-zero path wavedec2(level=0) degenerates to */
    int32_t reserved;
} WtLayerDesc;

typedef struct WtPlan WtPlan;

/* ---- metadata ------------------------------------------------------------------------------ */

/* library / ABI version (bumped on any signature change) */
int wt_version(void);
const char* wt_last_error(void);

/* pywt.Wavelet(name): id for the other calls and dec_len.  Names = the 17 of the --wavelet enum of
 * ResNet/main_pruning.py:89-96 (haar, db1/2/4/6, sym2/4/6, coif1/2/3, bior and rbio 1.3/2.2/4.4) plus db3/5/7/8/9, sym3,
 * bior1.1/3.1, rbio1.1/3.1 (dwt_pruning.py:35 takes any name).  Unknown name -> WT_E_INVALID (pywt raises ValueError). */
int wt_wavelet_lookup(const char* name, int* id, int* dec_len);
/* copies the four float64 filters (dec_lo, dec_hi, rec_lo, rec_hi), each `dec_len` long (HOST) */
int wt_wavelet_filters(int wavelet_id, double* dec_lo, double* dec_hi, double* rec_lo, double* rec_hi);

/* pywt.dwt_max_level(data_len, filter_len) as used by calculate_max_level, dwt_pruning.py:12-13 */
int wt_dwt_max_level(int data_len, int filter_len);
/* pywt.dwt_coeff_len(data_len, filter_len, mode) */
int wt_dwt_coeff_len(int data_len, int filter_len, int mode);
/* per-plane shape of pywt.coeffs_to_array(wavedec2(...))[0] (dwt_pruning.py:69-70) and the number
 * of real coefficients inside it (the rest is padding zeros that np.percentile counts) */
int wt_packed_shape(int H, int W, int wavelet_id, int level, int mode, int* side_h, int* side_w,
                    int64_t* coeffs_per_plane);

/* np.percentile's index arithmetic (method='linear') for a population of n values:
 * order-statistic ranks k0 <= k1 and interpolation weight gamma, under either NumPy semantic. */
int wt_percentile_plan(int64_t n, double percentile, int numpy_semantics, int64_t* k0, int64_t* k1,
                       double* gamma);

/* ---- the hot path: plans ------------------------------------------------------------------- */

/*
 * Build a plan that prunes `n_layers` tensors the way
 *   multi_resolution_analysis  (ResNet/dwt_pruning.py:35-95, threshold_kind PERCENTILE) or
 *   multi_resolution_analysis  (ResNet/dwt_pruning_NoEntropy.py:12-62, threshold_kind ABSOLUTE)
 * does for each of them: wavedec2(level=layers[i].level, mode, axes=(-2,-1)) -> coeffs_to_array ->
 * threshold -> array_to_coeffs -> waverec2 -> crop to the input shape -> zero count.
 * `threshold_value` is the percentile (0..100) or the absolute threshold.
 * host_io = 0: layers[i].in/out are device pointers (may alias).  host_io = 1: they are HOST
 * pointers; wt_plan_run then copies in, prunes and copies out (staging owned by the plan).
 * `layers` is copied; the caller may free it after the call.
 */
int wt_plan_create(const WtLayerDesc* layers, int n_layers, int wavelet_id, int mode,
                   int threshold_kind, double threshold_value, int numpy_semantics, int host_io,
                   WtPlan** plan_out);
int wt_plan_destroy(WtPlan* plan);
/* bytes of device memory the plan owns (coefficient workspace, histograms, tables, staging) */
size_t wt_plan_device_bytes(const WtPlan* plan);
/* number of kernel launches one wt_plan_run enqueues */
int wt_plan_num_launches(const WtPlan* plan);
/* re-point the plan at other tensors of identical shapes (same n_layers, batch, H, W, level) */
int wt_plan_set_io(WtPlan* plan, const WtLayerDesc* layers, int n_layers);
/* enqueue the whole pruning pass on `stream`; no host synchronisation for device-I/O plans */
int wt_plan_run(WtPlan* plan, void* stream);
/*
 * Results of the last run (HOST arrays; synchronises `stream`):
 *   counts[i*3+0] = numel               ("Original Parameter Count")
 *   counts[i*3+1] = nonzeros after      (dwt_pruning.py:119-120)
 *   counts[i*3+2] = zeros after         (dwt_pruning.py:88, torch.sum(pruned == 0))
 *   nonzero_before[i] = nonzeros of the input (dwt_pruning_NoEntropy.py:58,81); may be NULL
 *   thresholds[i] = the float32 value the strict `<` compared against; may be NULL
 */
int wt_plan_results(WtPlan* plan, void* stream, int64_t* counts, int64_t* nonzero_before,
                    float* thresholds);
/*
 * Optional device-side timing of the phases of wt_plan_run (memset, [h2d], one entry per analysis
 * launch group, select, one per synthesis launch group, [d2h]).  When enabled, wt_plan_run records
 * a cudaEvent on `stream` between phases; wt_plan_phase_times waits for the last run and returns
 * the number of phases (<= max_phases; negative = error), filling ms[i], launches[i] (kernels in
 * phase i) and the NUL-terminated names (names + i*name_stride).  Used by bench.py for the live
 * per-kernel durations; leave it off on the production path (each event costs ~1 us of stream time).
 */
int wt_plan_set_timing(WtPlan* plan, int enable);
/*
 * CUDA graphs.  enable = 1: the first two calls of wt_plan_run (one per alternating scratch block) capture the plan's
 * launch sequence -- every kernel of the run, the forks onto the plan's class streams and the joins back -- into a
 * cudaGraph; later runs replay it with ONE cudaGraphLaunch on `stream`.  Results are those of the plain run, bit for bit.
 * The reference has no counterpart (it calls PyWavelets / NumPy layer by layer: ResNet/dwt_pruning.py:67-77); this
 * is for callers that run one plan many times (threshold sweeps re-create plans and gain nothing).  The graphs are
 * dropped by anything that changes what a run launches (wt_plan_set_io / _set_peers / _set_multicast / _set_arithmetic)
 * and bypassed while phase timing is on.  A stream that cannot be captured (the legacy default stream) makes the plan
 * fall back to plain launches silently -- same results.  Device plans only run through it; a host-I/O plan forwards
 * the switch to the device plans of its layer groups.  Default (neither this call nor WT_GRAPH=0|1 in the environment when
 * the plan is created): automatic -- plans of four or more launches per run replay from their second run on (measured:
 * 1.5-2 % on conv-weight plans, 2-7 % on large-plane plans); single-launch plans (the level-0 flows) launch plainly.
 */
int wt_plan_set_graph(WtPlan* plan, int enable);
/* Arithmetic of the DWT multiply-adds.  WT_ARITH_FAITHFUL (default): fp32 multiply, then fp32 add, no FMA,
 * the order of oracle/dwt_oracle.c -- the arithmetic every bit-exactness claim of this library refers to.
 * WT_ARITH_FMA_TAPORDER: the same taps in the same order with every multiply-add fused into one FFMA (one rounding
 * instead of two).  A NAMED FAST MODE for the fp32-issue-bound decompositions (long filters on 3x3 planes), never chosen
 * silently: bit-exact against the oracle's twin (orc_set_arithmetic(1)), within 1e-5 relative / 1e-6 absolute of the
 * faithful weights, but exact zeros of a round trip are not preserved, so masks and zero counts can differ
 * from the reference's.  Covers the small-plane kernels and the tiled / large-plane kernels (every boundary mode). */
#define WT_ARITH_FAITHFUL 0
#define WT_ARITH_FMA_TAPORDER 1
int wt_plan_set_arithmetic(WtPlan* plan, int arithmetic);
/* which select the plan's percentile threshold runs through: 0 sampled (one full pass + short candidate lists),
 * 1 radix (three passes), 2 fused recompute kernels, 3 precompacted (the analysis kernels count and compact while they hold
 * the coefficients, the select only finishes the candidate lists), -1 none (absolute threshold / empty plan).  Reporting aid. */
int wt_plan_select_path(const WtPlan* plan);
/* Layer-sharded runs (north star: layers dealt to the GPUs of a box, the pruned state reassembled on every rank).
 * Every rank's output tensors live in a flat buffer of ONE layout that is mapped into all peers (CUDA VMM /
 * symmetric memory); byte_deltas[q] = (peer q's mapping of ITS buffer) - (this rank's buffer).  The kernels that
 * store pruned weights then store each value a second time at out + byte_deltas[q] for every q -- plain global
 * stores over NVLink, issued next to the arithmetic -- which replaces the all-gather after the last kernel.  The
 * caller still needs one cross-rank barrier before reading.  n_peers = 0 switches it off.  WT_E_UNSUPPORTED for plans
 * with layers on the generic tiled path or absolute thresholds (their kernels do not mirror stores). */
int wt_plan_set_peers(WtPlan* plan, int n_peers, const int64_t* byte_deltas);
/* The same over NVSwitch multicast (NVLS): byte_delta = (multicast mapping of the flat buffer, e.g. torch symmetric memory's
 * multicast_ptr) - (this rank's buffer).  Every pruned weight is then stored ONCE more, with multimem.st, and the switch
 * replicates it into every rank's buffer: a rank's NVLink egress is its shard instead of (ranks - 1) x its shard.
 * Percentile plans whose layers run on the small-plane or level-0 select kernels (WT_E_UNSUPPORTED otherwise). */
int wt_plan_set_multicast(WtPlan* plan, int64_t byte_delta);
int wt_plan_phase_times(WtPlan* plan, int max_phases, float* ms, int* launches, char* names,
                        int name_stride);
/* Staged runs for callers that take the order statistic outside the plan (no counterpart in the reference; used by the
 * intra-layer multi-GPU split).  stage 1 = every layer's analysis (coefficients stay in the plan's workspace),
 * stage 2 = threshold with the CURRENT contents of the plan's threshold array (wt_plan_thresholds_device: the caller writes
 * them between the stages) + synthesis + counts.  Device percentile plans on the spill kernels without peers
 * (WT_E_UNSUPPORTED otherwise); everything on `stream`, nothing synchronised.  wt_plan_results works after stage 2. */
int wt_plan_run_stage(WtPlan* plan, int stage, void* stream);
/* where stage 1 left layer `layer`'s coefficients: n_explicit floats at *coef (coefficient-major for conv-kernel planes,
 * band by band for large planes -- the ORDER is internal, the multiset is the packed array's nonzero cells and more), plus
 * n_implicit_zeros padding cells of pywt.coeffs_to_array that np.percentile also sees (dwt_pruning.py:70,27).  Level-0
 * layers: the input itself. */
int wt_plan_layer_coefficients(WtPlan* plan, int layer, const float** coef, int64_t* n_explicit, int64_t* n_implicit_zeros);
/* device address of the int64 [n_layers][4] result block of the LAST run: columns 1..3 = {nonzero, zeros,
 * nonzero_before} (column 0 is scratch: numel is known on the host).  Percentile plans alternate between two blocks,
 * so ask again after every wt_plan_run.  Valid in stream order after the run; nothing is synchronised. */
int wt_plan_counts_device(WtPlan* plan, const int64_t** counts_dev);
/* device address of the float32 [n_layers] thresholds the strict `<` compares against (percentile plans: written by
 * the run; absolute plans: the constant).  With wt_plan_counts_device this lets a caller sweep thresholds or chain
 * device-side logic without the stream synchronisation wt_plan_results implies. */
int wt_plan_thresholds_device(WtPlan* plan, const float** thresholds_dev);

/* one-shot convenience: create + run + results + destroy (device I/O). */
int wt_dwt_prune_f32(const WtLayerDesc* layers, int n_layers, int wavelet_id, int mode,
                     int threshold_kind, double threshold_value, int numpy_semantics, void* stream,
                     int64_t* counts, int64_t* nonzero_before, float* thresholds);

/* ---- building blocks (exposed for tests and for callers that keep coefficients) ------------- */
/* One radix pass of a DISTRIBUTED exact select (intra-layer multi-GPU split; the reference has no
 * counterpart: np.percentile sees one array, dwt_pruning.py:27).  hist[d] += number of elements of x whose key (the bit
 * pattern of |x|, 31 bits) has `prefix` in its top prefix_bits bits and d in the next digit_bits (<= 12) bits.  The caller
 * zeroes hist, sums it over the ranks (all-reduce) and descends: 12 + 10 + 9 bits give the exact k-th smallest |x| of the
 * union (wavelettransforms_b200/sharded.py: prune_tensor_split).  Device pointers; nothing is synchronised. */
int wt_key_hist_f32(const float* x, int64_t n, uint32_t prefix, int prefix_bits, int digit_bits, uint64_t* hist, void* stream);


/* Exact order statistics of |x| over a population of n explicit values plus `implicit_zeros`
 * zeros: ab_out[0] = sorted[k0], ab_out[1] = sorted[k1] (device float[2]).  Replaces the
 * np.partition inside np.percentile (dwt_pruning.py:27). */
int wt_select_abs_f32(const float* x, int64_t n, int64_t implicit_zeros, int64_t k0, int64_t k1,
                      float* ab_out, void* stream);

/* The same select for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * Its scratch (histograms, state, candidate keys: about 4.2 MB) is a fixed device array inside
 * the library that every GPU loading the library holds until exit, so the call allocates nothing; n is limited to
 * WT_SELECT_NOALLOC_MAX_N values (larger -> WT_E_INVALID, use wt_select_abs_f32).  All calls share the array: calls that
 * overlap on the GPU (different streams or host threads without ordering between them) race -- serialise them.
 * The results are written to ab_out in stream order; the call does not synchronise. */
#define WT_SELECT_NOALLOC_MAX_N (1 << 22)
int wt_select_abs_noalloc_f32(const float* x, int64_t n, int64_t implicit_zeros, int64_t k0, int64_t k1,
                              float* ab_out, void* stream);

/* wavedec2 + coeffs_to_array: x [batch][H][W] -> packed [batch][side_h][side_w] (padding = 0),
 * exactly the array dwt_pruning.py:69-70 hands to the thresholding. */
int wt_wavedec2_packed_f32(const float* x, int64_t batch, int H, int W, int wavelet_id, int level,
                           int mode, float* packed, void* stream);
/* array_to_coeffs + waverec2 + crop to [batch][H][W] (dwt_pruning.py:75-82) */
int wt_waverec2_packed_f32(const float* packed, int64_t batch, int H, int W, int wavelet_id,
                           int level, int mode, float* out, void* stream);

/* ---- sibling maskers ------------------------------------------------------------------------ */

/* percentage_min_pruning (ResNet/min_weight_pruning.py:66-74), in place, for a group of tensors:
 * zero the k[i] entries of smallest |w| of tensor i (ties: lowest flat index first), then
 * nonzero_out[i] = count_nonzero (min_weight_pruning.py:108-109).  w[i]: device pointers given in
 * a HOST array; n, k: HOST arrays; nonzero_out: HOST array (the call synchronises `stream`). */
int wt_min_weight_prune_f32(float* const* w, const int64_t* n, const int64_t* k, int n_tensors,
                            void* stream, int64_t* nonzero_out);

/* flat[idx[:k]] = 0 ; count_nonzero  (ResNet/random_pruning.py:53-58), in place.  idx: device
 * int64 indices produced by torch.randperm (kept as the RNG so masks match bit for bit).
 * nonzero_out: device int64[1]. */
int wt_scatter_zero_f32(float* w, int64_t n, const int64_t* idx, int64_t k, int64_t* nonzero_out,
                        void* stream);

/* count of x != 0 (torch.count_nonzero) and of |x| < eps (calculate_sparsity,
 * testing_suite/eval_model.py:7-20); out: device int64[2] = {nonzero, below_eps} */
int wt_count_f32(const float* x, int64_t n, float eps, int64_t* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WTPRUNE_H */
