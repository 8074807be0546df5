// Q: AWQ W4A16 (group 128) linear layers - placeholder translation unit, filled in below.
#include "t3b_common.cuh"

int32_t t3b_awq_gemv_launch(const void*, const t3b_linear_t*, int, void*, int, cudaStream_t) {
    T3B_FAIL(T3B_ERR_UNSUPPORTED, "AWQ GEMV not built yet");
}
int32_t t3b_awq_umma_launch(const void*, const t3b_linear_t*, int, void*, float*, int, cudaStream_t) {
    T3B_FAIL(T3B_ERR_UNSUPPORTED, "AWQ UMMA not built yet");
}
extern "C" int32_t t3b_repack_weights_awq(const int32_t*, const int32_t*, const void*, int32_t, int32_t, void*, void*, void*, void*) {
    T3B_FAIL(T3B_ERR_UNSUPPORTED, "AWQ repack not built yet");
}
