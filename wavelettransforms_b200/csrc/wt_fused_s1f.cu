// fused prune kernels for 1x1 planes, fma_taporder arithmetic -- see wt_fused_small.inc
#define WT_FS_H 1
#define WT_FS_W 1
#define WT_FS_FMA 1
#define WT_FS_NAME wt_fused_launch_s1f
#include "wt_fused_small.inc"
