// Layered tcgen05 path (placeholder until the general GEMM kernel lands).
#include "rfk_internal.cuh"

namespace rfk {

int tc_prepare(rfk_mlp* h) {
  (void)h;
  set_error("RFK_BF16 trials wider than 128 hidden units / 16 classes are not supported yet");
  return RFK_ERR_UNSUPPORTED;
}
void tc_release(rfk_mlp* h) { (void)h; }
size_t tc_workspace_bytes(const rfk_mlp* h, int64_t n) {
  (void)h;
  (void)n;
  return 0;
}
int tc_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes, cudaStream_t st) {
  (void)h; (void)x; (void)n; (void)probs; (void)ws; (void)ws_bytes; (void)st;
  set_error("layered tcgen05 path not available");
  return RFK_ERR_UNSUPPORTED;
}

}  // namespace rfk
