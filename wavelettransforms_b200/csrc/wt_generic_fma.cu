// wt_generic_fma.cu -- the generic / large-plane kernels in the named fast mode fma_taporder (include/wtprune.h:
// wt_plan_set_arithmetic): same kernels, same tap order, one FMA per tap instead of a multiply and an add.  Oracle twin:
// orc_set_arithmetic(1) (oracle/dwt_oracle.c).  Never the default.
#define WT_GENERIC_FMA 1
#include "wt_generic.cu"
