// fused prune kernels for 3x3 planes, faithful (no-FMA) arithmetic -- see wt_fused_small.inc
#define WT_FS_H 3
#define WT_FS_W 3
#define WT_FS_FMA 0
#define WT_FS_NAME wt_fused_launch_s3
#include "wt_fused_small.inc"
