// placeholder until the tcgen05 kernel lands
#include "common.cuh"
namespace b200 {
bool linear_tc_supported(int, int, int) { return false; }
int64_t linear_tc_workspace_bytes(int, int, int) { return 256; }
int linear_tc(const __nv_bfloat16*, int64_t, const __nv_bfloat16*, const __nv_bfloat16*, __nv_bfloat16*, int64_t, int,
              int, int, void*, int64_t, cudaStream_t) {
  set_error("tcgen05 linear not built");
  return B200_ERR_UNSUPPORTED;
}
}  // namespace b200
