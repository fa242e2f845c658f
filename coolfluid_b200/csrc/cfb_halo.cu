// cfb200 — multi-GPU entry points.  Placeholder translation unit: filled in by the
// halo-exchange milestone; until then the calls report CFB200_ERR_UNSUPPORTED.
#include "../../include/cfb200.h"
extern "C" {
int cfb200_nccl_unique_id(void*) { return CFB200_ERR_UNSUPPORTED; }
int cfb200_comm_init(cfb200_handle*, int, int, const void*) { return CFB200_ERR_UNSUPPORTED; }
int cfb200_set_halo(cfb200_handle*, const int32_t*, const int32_t*, const int32_t*, const int32_t*) { return CFB200_ERR_UNSUPPORTED; }
int cfb200_halo_exchange(cfb200_handle*) { return CFB200_ERR_UNSUPPORTED; }
int cfb200_allreduce_sum(cfb200_handle*, double*, int) { return CFB200_ERR_UNSUPPORTED; }
}
