// gemm_tcgen05.cu -- prefill / batched-decode regime (M >= 64): tcgen05 kind::i8 GEMM.
#include "common.cuh"

namespace b200 {

bool gemm_i8_supported(const bitnet_b200_weight*, int) { return false; }

int launch_gemm_i8(const bitnet_b200_ctx*, const bitnet_b200_weight*, const int8_t*, const float*,
                   int, float*, int32_t*) {
    return fail("tcgen05 GEMM not built");
}

}  // namespace b200
