// wt_fused_flat.cu -- the flat (level-0) class of the fused prune kernel, the chunk geometry the plan builds its ticket
// queue from, and the class dispatcher.
#include "wt_fused_kinds.cuh"

int wt_fused_launch_s1(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s3(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s5(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s7(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s1f(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s3f(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s5f(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);
int wt_fused_launch_s7f(const WtBank&, const WtFArgs&, int, double, cudaStream_t, int*);

bool wt_fused_supported(int H, int W, int F, int level, int mode) {
    if (level == 0) return true;  // flat class: any shape, any mode (wavedec2(level=0) is the identity)
    if (mode != WT_MODE_PERIODIZATION || level > WT_SMALL_MAX_LEVEL) return false;
    if (H != W || !(H == 1 || H == 3 || H == 5 || H == 7)) return false;
    return F == 2 || F == 4 || F == 6 || F == 8 || F == 10 || F == 12 || F == 18;
}

// S samples a job only while the sample keeps 16 Ki keys
static int sample_shift(long long keys) {
    int s = 0;
    while (s < 4 && (keys >> (s + 1)) >= (1LL << 14)) ++s;
    return s;
}

void wt_fused_geometry(int H, int W, WtFJob* job, long long keys_per_unit) {
    const long long units = job->units;
    job->n = units * keys_per_unit;
    job->sample_shift = sample_shift(job->n);
    const int s = job->sample_shift;
    if (H == 0) {  // flat: items of WT_F_FLAT_CHUNK elements; a sample item covers 16 tiles of 4096 elements
        job->items = (int)((units + WT_F_FLAT_CHUNK - 1) / WT_F_FLAT_CHUNK);
        const long long tiles = (units + 4095) >> 12;
        job->s_items = s == 0 ? job->items : (int)((tiles + 15) / 16);
    } else {       // small planes: one item = one tile of planes; a sample item = one tile out of every 2^s
        const int tile = WT_F_SMALL_THREADS(H, W);
        const long long tiles = (units + tile - 1) / tile;
        job->items = (int)tiles;
        job->s_items = (int)((tiles + (1LL << s) - 1) >> s);
    }
}

int wt_fused_launch(int H, int W, const WtBank& bank, int fma, const WtFArgs& args, int reserve, double share, cudaStream_t stream,
                    int* grid_out) {
    if (grid_out) *grid_out = 0;
    if (args.total_items <= 0) return WT_OK;
    if (H == 0 && W == 0) return wtf::launch<wtf::FlatKind>(bank, args, reserve, share, stream, grid_out);
#ifdef WT_WITH_FMA_KERNELS
#define WT_PICK(n) (fma ? wt_fused_launch_s##n##f(bank, args, reserve, share, stream, grid_out) : wt_fused_launch_s##n(bank, args, reserve, share, stream, grid_out))
#else
    WT_CHECK(!fma, WT_E_UNSUPPORTED, "this build of libwtprune has no fma_taporder kernels");
#define WT_PICK(n) wt_fused_launch_s##n(bank, args, reserve, share, stream, grid_out)
#endif
    if (H == 1 && W == 1) return WT_PICK(1);
    if (H == 3 && W == 3) return WT_PICK(3);
    if (H == 5 && W == 5) return WT_PICK(5);
    if (H == 7 && W == 7) return WT_PICK(7);
#undef WT_PICK
    wt_set_error("fused path: unsupported plane %dx%d", H, W);
    return WT_E_UNSUPPORTED;
}
