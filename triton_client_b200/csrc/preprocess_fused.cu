// Fused single-kernel preprocess path (placeholder until the kernel lands: nothing is eligible,
// every image takes the two-pass kernels of preprocess.cu).
#include "preprocess.cuh"

namespace tcb {

bool fused_supported(const Plan&, const ImgDesc&, int*, size_t*) { return false; }

int fused_launch(const Plan&, void*, cudaStream_t) {
  return set_error(TCB_ERR_INTERNAL, "fused preprocess kernel not built");
}

}  // namespace tcb
