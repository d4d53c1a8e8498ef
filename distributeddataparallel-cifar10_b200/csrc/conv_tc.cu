// placeholder until the tcgen05 kernels land
#include "nrd_common.cuh"
int nrd_tc_plan_init(nrd_plan* p) { p->tc_ready = 0; return NRD_OK; }
int nrd_tc_conv_fwd(nrd_plan*, int, cudaStream_t) { nrd_set_error("tcgen05 path not built"); return NRD_EUNSUPPORTED; }
int nrd_tc_conv_dgrad(nrd_plan*, const void*, void*, cudaStream_t) { nrd_set_error("tcgen05 path not built"); return NRD_EUNSUPPORTED; }
int nrd_tc_conv_wgrad(nrd_plan*, int, int, cudaStream_t) { nrd_set_error("tcgen05 path not built"); return NRD_EUNSUPPORTED; }
int nrd_tc_stats_blocks(const nrd_plan*) { return 0; }
