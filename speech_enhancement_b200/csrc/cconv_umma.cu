// tcgen05 bf16x3 implicit-GEMM engine for the complex convolutions (placeholder until the
// kernel lands: reports "unsupported" so SE_ENGINE_AUTO stays on the SIMT engine).
#include "engines.cuh"
namespace se {
bool cconv_umma_supported(bool, int64_t, const ConvGeom&) { return false; }
int64_t cconv_umma_workspace_bytes(bool, int64_t, int, int, int, int, int, int, int, int) { return 0; }
int cconv_fwd_umma(bool, const float*, const float*, const float*, int64_t, const ConvGeom&, float*,
                   void*, int64_t, cudaStream_t) {
  return fail(SE_ERR_UNSUPPORTED, "se_cconv*_fwd: tcgen05 engine not built");
}
}  // namespace se
