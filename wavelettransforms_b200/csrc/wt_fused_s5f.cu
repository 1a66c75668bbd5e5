// fused prune kernels for 5x5 planes, fma_taporder arithmetic -- see wt_fused_small.inc
#define WT_FS_H 5
#define WT_FS_W 5
#define WT_FS_FMA 1
#define WT_FS_NAME wt_fused_launch_s5f
#include "wt_fused_small.inc"
