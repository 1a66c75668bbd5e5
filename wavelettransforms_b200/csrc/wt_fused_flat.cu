// wt_fused_flat.cu -- the flat (level-0) class of the fused prune kernel, the chunk geometry the plan builds its ticket
// queue from, and the class dispatcher.
#include "wt_fused_kinds.cuh"

int wt_fused_launch_s1(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s3(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s5(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s7(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s1f(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s3f(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s5f(const WtBank&, const WtFArgs&, cudaStream_t);
int wt_fused_launch_s7f(const WtBank&, const WtFArgs&, cudaStream_t);

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
    if (H == 0) {  // flat: items of WT_F_FLAT_CHUNK elements; sample items of 2 << s tiles of 4096
        job->chunks = (int)((units + WT_F_FLAT_CHUNK - 1) / WT_F_FLAT_CHUNK);
        const long long tiles = (units + 4095) >> 12;
        job->s_chunks = s == 0 ? job->chunks : (int)((tiles + (2LL << s) - 1) / (2LL << s));
    } else {       // small planes: items of WT_F_SMALL_CHUNK_TILES tiles; S takes one tile out of every 2^s
        const int tile = WT_F_SMALL_THREADS(H, W);
        const long long tiles = (units + tile - 1) / tile;
        job->chunks = (int)((tiles + WT_F_SMALL_CHUNK_TILES - 1) / WT_F_SMALL_CHUNK_TILES);
        const long long groups = (tiles + (1LL << s) - 1) >> s;
        job->s_chunks = (int)((groups + WT_F_SMALL_CHUNK_TILES - 1) / WT_F_SMALL_CHUNK_TILES);
    }
    long long f = job->n >> 16;
    job->f_chunks = (int)(f < 1 ? 1 : (f > 32 ? 32 : f));
}

int wt_fused_launch(int H, int W, const WtBank& bank, int fma, const WtFArgs& args, cudaStream_t stream) {
    if (args.n_tickets <= 0) return WT_OK;
    if (H == 0 && W == 0) return wtf::launch<wtf::FlatKind>(bank, args, 0, stream);
    if (H == 1 && W == 1) return fma ? wt_fused_launch_s1f(bank, args, stream) : wt_fused_launch_s1(bank, args, stream);
    if (H == 3 && W == 3) return fma ? wt_fused_launch_s3f(bank, args, stream) : wt_fused_launch_s3(bank, args, stream);
    if (H == 5 && W == 5) return fma ? wt_fused_launch_s5f(bank, args, stream) : wt_fused_launch_s5(bank, args, stream);
    if (H == 7 && W == 7) return fma ? wt_fused_launch_s7f(bank, args, stream) : wt_fused_launch_s7(bank, args, stream);
    wt_set_error("fused path: unsupported plane %dx%d", H, W);
    return WT_E_UNSUPPORTED;
}
