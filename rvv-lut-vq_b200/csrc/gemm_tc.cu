// gemm_tc.cu -- placeholder until the tcgen05 prefill kernel lands.
#include "common.cuh"
namespace lutvq {
size_t gemm_workspace_bytes(int, int) { return 0; }
int gemm_dequant_tc_launch(const void*, const void*, const void*, const void*, const void*, void*, int, int, int, int,
                           void*, size_t, cudaStream_t) {
    return fail(LUTVQ_EUNSUPPORTED, "tcgen05 prefill GEMM not built yet");
}
}  // namespace lutvq
