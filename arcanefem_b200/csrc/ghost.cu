// ghost.cu — multi-GPU ghost-row exchange (placeholder until the NCCL path lands).
#include "internal.h"

namespace b2f {
int ghost_plan(b2f_context_s*, int, int, int64_t, const int32_t*, const int32_t*, const int64_t*)
{
  return fail(B2F_ERR_NOT_IMPLEMENTED, "b2f_ghost_plan: not built yet");
}
int ghost_exchange(b2f_context_s*, void*, cudaStream_t)
{
  return fail(B2F_ERR_NOT_IMPLEMENTED, "b2f_ghost_exchange: not built yet");
}
} // namespace b2f

extern "C" {
int b2f_nccl_unique_id(void*) { return b2f::fail(B2F_ERR_NOT_IMPLEMENTED, "b2f_nccl_unique_id: not built yet"); }
int b2f_nccl_comm_init(const void*, int32_t, int32_t, void**) { return b2f::fail(B2F_ERR_NOT_IMPLEMENTED, "b2f_nccl_comm_init: not built yet"); }
int b2f_nccl_comm_destroy(void*) { return b2f::fail(B2F_ERR_NOT_IMPLEMENTED, "b2f_nccl_comm_destroy: not built yet"); }
}
