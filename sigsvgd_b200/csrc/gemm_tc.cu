// placeholder until the tcgen05 kernels land: the SIMT fp32 path handles every size.
#include "gemm_tc.cuh"

namespace sigsvgd {

size_t tc_workspace_bytes(int64_t, int64_t, int64_t, int64_t) { return 0; }
bool tc_path_enabled(int64_t, int64_t, int64_t, int64_t, uint32_t) { return false; }
int tc_pack_and_sqdist(const float*, int64_t, int64_t, int64_t, int64_t, float*, void*, cudaStream_t) {
  set_error("tcgen05 path not built");
  return SIGSVGD_ERR_UNSUPPORTED;
}
int tc_matmul_kv(const float*, int64_t, const float*, int64_t, int64_t, int64_t, float*, int64_t,
                 void*, cudaStream_t) {
  set_error("tcgen05 path not built");
  return SIGSVGD_ERR_UNSUPPORTED;
}

}  // namespace sigsvgd
