// wt_common.cuh -- shared declarations of libwtprune (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <vector>

#include "../../include/wtprune.h"

#define WT_MAX_FILTER 18
#define WT_MAX_LEVEL 32

void wt_set_error(const char* fmt, ...);

#define WT_CUDA(expr)                                                                      \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            wt_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                         __LINE__);                                                        \
            return WT_E_CUDA;                                                              \
        }                                                                                  \
    } while (0)

#define WT_CHECK(cond, code, ...)      \
    do {                               \
        if (!(cond)) {                 \
            wt_set_error(__VA_ARGS__); \
            return (code);             \
        }                              \
    } while (0)

// ---- device info ------------------------------------------------------------------------------
struct WtDeviceInfo {
    int device;
    int sm_count;
    int coop;
};
int wt_device_info(WtDeviceInfo* info);

// ---- wavelet banks (fp32 casts of the generated double tables) ---------------------------------
struct WtBank {
    int F;
    float dec_lo[WT_MAX_FILTER], dec_hi[WT_MAX_FILTER], rec_lo[WT_MAX_FILTER], rec_hi[WT_MAX_FILTER];
};
int wt_get_bank(int wavelet_id, WtBank* bank);
int wt_idwt_len(int n, int F, int mode);
// sizes[0] = n, sizes[l] = coefficient length at level l
void wt_level_sizes(int n, int F, int mode, int level, int* sizes);

// Layer-sharded runs (sharded.py): every rank owns a flat buffer of the same layout, mapped into every peer (CUDA VMM /
// symmetric memory); a kernel that stores a pruned weight also stores it at the same offset of each peer's buffer --
// plain st.global over NVLink -- so the all-gather happens under the arithmetic instead of after it.
#define WT_MAX_PEERS 7
struct WtPeers {
    int n;
    int mc;                         // 1: delta[0] is the distance to the MULTICAST mapping of the buffer (NVLS): one
                                    // multimem.st per value, replicated to every rank inside the NVSwitch
    long long delta[WT_MAX_PEERS];  // byte distance from this rank's buffer to the peer's mapping of its own
};
#ifdef __CUDACC__
// stores through a multicast (NVLS) address: the switch replicates the value into every GPU of the multicast group
__device__ __forceinline__ void wt_mc_store4(void* p, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void wt_mc_store1(void* p, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
#endif

// ---- select (wt_select.cu) ----------------------------------------------------------------------
// One job = one population whose order statistics are wanted: the explicit values x[0..n) plus
// `zeros` implicit zeros (coeffs_to_array padding).  Jobs with out != nullptr also get the
// level-0 apply phase (out = |x| < t ? 0 : x) and zero counting fused into the same launch.
#define WT_SEL_TILE 4096
#define WT_SEL_THREADS 512
#define WT_SEL_BINS1 4096  // key bits 30..19
#define WT_SEL_BINS2 1024  // key bits 18..9
#define WT_SEL_BINS3 512   // key bits  8..0
#define WT_SEL_COARSE (WT_SEL_BINS1 / 32)
#define WT_SEL_CBINS1 4096 // candidate pass C1 of the sampled path: top 12 bits of key - lo_key inside the bracket
#define WT_SEL_CBINS 2048  // candidate pass C2: the next <= 11 bits, one histogram per rank
#define WT_SEL_CBINS3 256  // candidate pass C3 (brackets wider than 16 bins only): the last <= 8 bits

enum { WT_LERP_LEGACY = 0, WT_LERP_NEP50 = 1, WT_LERP_ABSOLUTE = 2 };

struct WtSelJob {
    const float* x;
    float* out;          // nullable
    unsigned int* cand;  // candidate keys of the fast path (nullable: then the job takes the 3-pass path)
    long long n;
    long long zeros;
    long long k0, k1;
    double gamma;
    float abs_threshold;
    int lerp_mode;
    int tile_begin;  // first tile of this job in the flattened tile space
    int n_tiles;
    int layer;       // index into the counts / thresholds arrays
    unsigned int cand_cap;  // capacity of cand[] in keys
    int sample_shift;       // sampled path: S0 reads 4096 >> sample_shift elements of every 4096-element tile
    WtPeers peers;          // level-0 apply: out is also stored at out + delta of every peer (read from the job table)
};

struct WtSelState {  // per job, device; written by the CTA that holds the job's first tile
    unsigned int prefix1[2];       // 12-bit digits of rank k0 / k1 (after P1)
    unsigned int prefix2[2];       // 22-bit prefixes (after P2)
    unsigned long long rem1[2];    // ranks inside those digits
    unsigned long long rem2[2];
    float a, b, t;                 // sorted[k0], sorted[k1], threshold
    unsigned int cand_count;       // radix path: keys appended to cand[] by P2
    int overflow;                  // 1: more candidates than cand_cap -> P3 re-reads the tensor
    int failed;                    // sampled path gave up on this job (bracket missed / list overflow): radix path runs
    // sampled path
    unsigned long long below;      // explicit keys < lo_bin << 19 (counted by S1)
    unsigned int s_count;          // keys appended to cand[] by S1 (keys inside the bracket)
    unsigned int lo_bin, n_bins;   // bracket [lo_bin, lo_bin + n_bins) in 12-bit bins; n_bins == 0: S1 did not compact
    unsigned int cprefix[2];       // C1 digit of both ranks (published by C2)
    unsigned long long crem[2];    // ranks inside those digits
    unsigned int cprefix2[2];      // wide brackets: C1 and C2 digits of both ranks (published by C3)
    unsigned long long crem2[2];
    unsigned int tail_key[2];      // one-CTA tail: key - lo_key of both ranks
    int tail_done;                 // 1: the tail finished this job (C1..C3 skip it, final_keys reads tail_key)
};

struct WtSelArgs {
    const WtSelJob* jobs;
    int n_jobs;
    int n_tiles;
    WtSelState* state;
    unsigned long long* hist1;  // [n_jobs][BINS1]
    unsigned long long* hist2;  // [n_jobs][CBINS1]  (radix path: [2][BINS2] inside the same slot)
    unsigned long long* hist3;  // [n_jobs][2][CBINS]  (radix path: [2][BINS3] inside the same slot)
    int* fail_flag;             // set when any job leaves the sampled path
    int* wide_flag;             // set when any job's bracket is wider than 16 bins: the grid runs candidate pass C3
    unsigned int tail_max;      // longest candidate list the one-CTA tail takes (resolved at launch; 0: no tail)
    int short_lists;            // host hint (wt_select_predict_short): every job's candidate list should fit the one-CTA tail
    int* big_flag;              // set when any job's candidate list is too long for the one-CTA tail: the grid runs C1..C3
    int path;                   // 0: sampled path with radix fallback, 1: radix path only; set by wt_select_launch
    unsigned long long* hist1c; // [n_jobs][BINS1 / 32]: coarse sums of hist1 (32 fine bins each)
    const int* cta_tiles;       // nullable: [cta_grid + 1] first tile of every CTA (cost-balanced split)
    const int* cta_job;         // nullable: [cta_grid] job of that first tile
    int cta_grid;               // grid size the two tables were built for
    unsigned long long* stamps; // nullable tuning aid: [grid][16] per-CTA phase stamps / accumulated ns
    float* thresholds;          // [n_layers]
    long long* counts;          // [n_layers][4] {numel, nonzero, zeros, nonzero_before}
    int any_apply;              // at least one job has out != nullptr
    int precompacted;           // 1: the small-plane analysis kernels have done S0 / S1 for the jobs with sample_shift >= 0 (sample
                                // histogram -> wt_select_bracket_launch -> count + compact while the coefficients were in registers):
                                // the select starts at the tail; jobs with sample_shift < 0 (level 0, generic path) take the radix passes
    char* zero_next;            // nullable: a block this launch zeroes for the NEXT run of the plan (two scratch blocks alternate,
    unsigned long long zero_next_bytes;  // so a step enqueues no memset); multiple of 16 bytes
};

// candidate capacity (keys) the plan reserves for a population of n explicit values
static inline unsigned int wt_select_cand_cap(long long n) {
    long long c = n / 4;
    if (c > (1LL << 21)) c = 1LL << 21;
    if (c < 4096) c = 4096;
    if (c > n) c = n;
    return (unsigned int)c;
}

// S0 sample of the sampled path: at least 64 Ki keys, 1/16 of the tensor from 1 Mi keys on
#ifndef WT_SEL_MIN_SAMPLE_LOG2
#define WT_SEL_MIN_SAMPLE_LOG2 14  // a job is sub-sampled only while the sample keeps 16 Ki keys
#endif
static inline int wt_select_sample_shift(long long n) {
    static const int min_log2 = [] { const char* e = getenv("WT_SEL_MIN_SAMPLE_LOG2"); return e ? atoi(e) : WT_SEL_MIN_SAMPLE_LOG2; }();
    int s = 0;
    while (s < 4 && (n >> (s + 1)) >= (1LL << min_log2)) ++s;
    return s;
}
size_t wt_select_scratch_bytes(int n_jobs);  // hist1+hist2+hist3+done+state, contiguous, memset to 0 before launch
// lays out args over a zeroed scratch block
void wt_select_bind(WtSelArgs* a, void* scratch, int n_jobs);
// only re-points the scratch-relative fields (histograms, state, flags) at another zeroed block of the same layout
void wt_select_rebind(WtSelArgs* a, void* scratch, int n_jobs);
int wt_select_launch(const WtSelArgs& args, cudaStream_t stream);
// precompacted launches: one CTA per job turns the sample histogram (hist1 / hist1c) into the bracket of 12-bit bins the
// analysis kernels compact (state.lo_bin / n_bins, 0 bins = none); jobs with sample_shift < 0 are handed to the radix passes
int wt_select_bracket_launch(const WtSelArgs& args, cudaStream_t stream);
#define WT_SEL_SAMPLED_MIN_TILES 5500   // 22.5 M keys = 90 MB: beyond that the keys are not L2-resident between the radix path's two passes
#ifndef WT_SEL_TAIL_MAX
#define WT_SEL_TAIL_MAX 32768  // longest candidate list one CTA finishes alone (three L2-resident passes; measured
                               // against the grid-wide passes: wins at 20 K keys, loses at 50 K)
#endif
int wt_select_path(int n_tiles, int short_lists);  // 0 sampled, 1 radix
// plan-time guess: will every job's bracket hold few enough keys for the one-CTA tail? (see wt_select.cu)
int wt_select_predict_short(const WtSelJob* jobs, int n_jobs);
// grid wt_select_launch will use for n_tiles tiles on the current device
int wt_select_grid(int n_tiles, int* grid_out);
// cost-balanced contiguous split of the tile space over `grid` CTAs: every segment (piece of one job inside one
// CTA) costs WT_SEL_SEGMENT_COST tiles on top of its tiles.  tiles_out[grid + 1], job_out[grid] (HOST)
#define WT_SEL_SEGMENT_COST 12
void wt_select_balance(const WtSelJob* jobs, int n_jobs, int n_tiles, int grid, int* tiles_out, int* job_out);
#define WT_SEL_STAMP_CTAS 2048

// ---- fused prune kernels (wt_fused.cuh): one cooperative launch per plane class, coefficients never leave the SM ------
#define WT_F_BINS1 4096   // key bits 30..19
#define WT_F_FBINS 2048   // 11-bit refinement digits
#define WT_F_FBINS3 512   // last <= 9 bits
#define WT_F_HIST_WORDS (WT_F_BINS1 + WT_F_BINS1 / 32 + WT_F_FBINS + 2 * WT_F_FBINS + 2 * WT_F_FBINS3)  // u32 per job
#define WT_F_FLAT_CHUNK 4096                                        // flat class: elements per item
#define WT_F_SMALL_THREADS(H, W) (((H) * (W) <= 9) ? 256 : 128)     // small-plane classes: planes per item = threads per CTA
#define WT_F_WFLUSH 224   // a warp moves its staged candidates to the global list once it holds more keys
#define WT_F_GROUP 8      // keys per thread the count / compact pass handles at once
#define WT_F_WSTAGE (WT_F_WFLUSH + 32 * WT_F_GROUP)   // one more group always fits
#define WT_F_NOKEY 0xffffffffu   // not a key (ragged end of an item, lane without a plane): beyond every bracket
#define WT_F_JIDX 512     // jobs whose item offsets a CTA keeps in shared memory (more: looked up in global memory)
#define WT_F_SLICE 4096   // candidate keys per slice of the tail rounds
#define WT_F_STAMP_SLOTS 16      // u64 slots per CTA of the tuning stamps (WT_FUSED_STAMPS)

struct __align__(16) WtFJob {   // one tensor of the class (device, read-only during a run)
    const float* in;         // [units] elements (flat class) or [units][H][W] planes
    float* out;
    unsigned int* cand;      // candidate keys (nullable: the job then takes the exact path X)
    long long units;         // flat class: elements; small-plane classes: planes
    long long n;             // explicit keys of the population (flat: units; small: units * coefficients per plane)
    long long zeros;         // implicit zeros (coeffs_to_array padding)
    long long k0, k1;        // np.percentile's order-statistic ranks in the population of n + zeros
    double gamma;
    float abs_threshold;
    int lerp_mode;           // WT_LERP_LEGACY / WT_LERP_NEP50 / WT_LERP_ABSOLUTE (no select: apply only)
    unsigned int cand_cap;
    int level;
    int layer;               // row of counts[] / thresholds[]
    int sample_shift;        // S reads 1 / 2^shift of the job
    int items, s_items;      // items of the full passes / of the sample pass
    int item_begin, s_begin; // first item of this job in the class-wide item spaces (prefix sums over the jobs)
    WtPeers peers;           // layer-sharded runs: out is also stored at out + delta of every peer
};
static_assert(sizeof(WtFJob) % 16 == 0, "WtFJob is copied with 16-byte loads");

struct __align__(128) WtFState {   // per job, inside the zeroed scratch; one 128-byte line per access pattern
    // line 0: written once per phase by the CTA that holds the job's first item / slice, read in the next phase
    unsigned long long r[2], r2[2];   // ranks inside the digits found so far
    unsigned int lo_key, width;       // bracket [lo_key, lo_key + width) in key patterns (0 width: no bracket)
    unsigned int d[2], d2[2];         // digits of both ranks after the F1 / F2 (X1 / X2) scans
    int failed;                       // the job left the sampled pipeline: the exact path X runs it
    float a, b, t;                    // sorted[k0], sorted[k1], threshold
    int pad0[10];
    // line 1: the candidate list's fill level (an atomicAdd with a consumer inside the C pass)
    unsigned int s_count;             // keys appended to cand[] by C
    int pad1[31];
    // line 2: accumulators and the completion counter of the apply (fire-and-forget atomics, one consumer each)
    unsigned long long below;         // explicit keys < lo_key (counted by C)
    unsigned long long zc, nzb;       // zeros written / nonzero inputs (A)
    unsigned int done;                // items of the apply pass completed
    int pad2[25];
};
static_assert(sizeof(WtFState) == 384, "WtFState: three 128-byte lines");

struct WtFHeader {           // first 256 bytes of the scratch
    unsigned int fail_count; // jobs handed to the exact path
    unsigned int wide;       // some bracket is wider than 8 bins: the tail runs its second round
    int pad[62];
};

struct WtFArgs {
    const WtFJob* jobs;
    int n_jobs;
    int n_select;            // jobs that run a select (0: absolute-threshold class, phase A only, no grid barrier)
    int total_items, total_s_items;
    char* scratch;           // this run's block: zero at launch
    char* scratch_next;      // zeroed by this run for the next one
    unsigned long long scratch_bytes;
    float* thresholds;       // [n_layers]
    long long* counts;       // [n_layers][4] {numel, nonzero, zeros, nonzero_before}
    unsigned long long* stamps;  // nullable tuning aid: [grid][WT_F_STAMP_SLOTS] phase boundaries per CTA (WT_FUSED_STAMPS)
};

static inline size_t wt_fused_scratch_bytes(int n_jobs) {
    size_t b = sizeof(WtFHeader) + (size_t)n_jobs * sizeof(WtFState) + (size_t)n_jobs * WT_F_HIST_WORDS * 4;
    return (b + 255) / 256 * 256;
}
// item geometry of a job (fills n, sample_shift, items, s_items from units); H = W = 0: flat class
void wt_fused_geometry(int H, int W, WtFJob* job, long long keys_per_unit);
// true when (H, W, F, level) has a fused instantiation
bool wt_fused_supported(int H, int W, int F, int level, int mode);
// Cooperative launch of the class's kernel (H = W = 0: flat); fma: arithmetic fma_taporder.  The grid is what fits the
// device minus `reserve_ctas` (room for the kernels of other classes that run next to it), at most one CTA per item and
// at most `share` (0..1: this class's part of the plan's bytes) of the device; *grid_out = the grid used.
int wt_fused_launch(int H, int W, const WtBank& bank, int fma, const WtFArgs& args, int reserve_ctas, double share,
                    cudaStream_t stream, int* grid_out);

// ---- level-0 apply for absolute thresholds and generic elementwise helpers (wt_select.cu) --------
int wt_apply_threshold_launch(const float* x, float* out, long long n, const float* thr_dev, float thr_host,
                              long long* counts4, cudaStream_t stream);

// ---- small-plane regime (wt_small.cu): one thread per kH x kW plane, all levels in registers -----
#define WT_SMALL_THREADS 128
#define WT_SMALL_MAX_LEVEL 6

struct WtSmallJob {
    const float* in;   // [batch][H][W]
    float* out;        // [batch][H][W]
    float* coef;       // [ncoef][batch]  (coefficient-major so a warp's accesses coalesce)
    long long batch;
    int level;
    int layer;
    int tile_begin;
    int n_tiles;
    float abs_threshold;
    int use_abs;       // 1: compare against abs_threshold, 0: against thresholds[layer]
    // precompacted select (percentile plans): the analysis kernel counts the keys below the job's bracket and appends the keys
    // inside it to the select job's candidate list while they are in registers
    int sel_job;       // index of the tensor's select job (-1: none)
    int sample_shift;  // the sample pass analyses one tile out of every 2^shift
    unsigned int* cand;
    unsigned int cand_cap;
    int pad;
    WtPeers peers;     // synthesis: the output tile is also stored at out + delta of every peer
};

// true when (H, W, F, level, mode) has a register-resident instantiation
bool wt_small_supported(int H, int W, int F, int level, int mode);
// number of real coefficients per plane (periodization)
int wt_small_ncoef(int H, int W, int level);
// fma: 0 faithful arithmetic (multiply, then add: the oracle's and the reference's), 1 fma_taporder (same tap order, fused)
int wt_small_analysis_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                             long long* counts, cudaStream_t stream);
// precompacted select: (1) sample pass -- 12-bit histogram of |c| over one tile in 2^shift of every job, into the select's
// hist1 / hist1c of the job; (2) after wt_select_bracket_launch: the analysis with count + compact against state[sel_job]
int wt_small_sample_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                           unsigned long long* hist1, unsigned long long* hist1c, cudaStream_t stream);
int wt_small_analysis_compact_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                                     long long* counts, WtSelState* state, cudaStream_t stream);
int wt_small_synthesis_launch(int H, int W, const WtBank& bank, int fma, const WtSmallJob* jobs_dev, int n_jobs, int n_tiles,
                              const float* thresholds, long long* counts, cudaStream_t stream);
