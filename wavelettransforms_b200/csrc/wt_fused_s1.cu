// fused prune kernels for 1x1 planes, faithful (no-FMA) arithmetic -- see wt_fused_small.inc
#define WT_FS_H 1
#define WT_FS_W 1
#define WT_FS_FMA 0
#define WT_FS_NAME wt_fused_launch_s1
#include "wt_fused_small.inc"
