// fused prune kernels for 7x7 planes, faithful (no-FMA) arithmetic -- see wt_fused_small.inc
#define WT_FS_H 7
#define WT_FS_W 7
#define WT_FS_FMA 0
#define WT_FS_NAME wt_fused_launch_s7
#include "wt_fused_small.inc"
