// fused prune kernels for 3x3 planes, fma_taporder arithmetic -- see wt_fused_small.inc
#define WT_FS_H 3
#define WT_FS_W 3
#define WT_FS_FMA 1
#define WT_FS_NAME wt_fused_launch_s3f
#include "wt_fused_small.inc"
