// the small-plane analysis / synthesis kernels in the `fma_taporder` arithmetic -- see wt_small.cu
#define WT_SMALL_FMA 1
#include "wt_small.cu"
