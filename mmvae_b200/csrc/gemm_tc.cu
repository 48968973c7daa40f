// Kernel (4), tensor-core path: tcgen05.mma (kind::tf32, 3xTF32 error-compensated split) fed by TMA, accumulators
// in TMEM.  Placeholder until the kernel lands: reports "unsupported" so the dispatcher uses the exact-fp32 SIMT GEMM.
#include "common.cuh"

int mmvae_gemm_tc_supported(int, int, int, int, int) { return 0; }
int mmvae_gemm_tc(const float*, const float*, const float*, float*, int, int, int, int, int, int, float*, long,
                  cudaStream_t) {
    return mmvae_set_error(1, "gemm_tc: not built");
}
extern "C" int mmvae_has_tcgen05(void) { return 0; }
