// fused prune kernels for 5x5 planes, faithful (no-FMA) arithmetic -- see wt_fused_small.inc
#define WT_FS_H 5
#define WT_FS_W 5
#define WT_FS_FMA 0
#define WT_FS_NAME wt_fused_launch_s5
#include "wt_fused_small.inc"
