// Internal entry points shared between the translation units of librafiki_b200.
#pragma once
#include "rfk_common.cuh"

namespace rfk {

// mlp_f32.cu -- fp32 CUDA-core path
size_t mlp_f32_workspace_bytes(const rfk_mlp* h, int64_t n);
int mlp_f32_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes,
                    cudaStream_t st);

// gemm_tc.cu -- layered tcgen05 path (decode -> GEMM+bias+ReLU ... -> GEMM+bias+softmax)
int tc_prepare(rfk_mlp* h);
void tc_release(rfk_mlp* h);
size_t tc_workspace_bytes(const rfk_mlp* h, int64_t n);
int tc_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes,
               cudaStream_t st);

// fused_trial.cu -- single-kernel tcgen05 path for TfFeedForward-shaped trials
bool fused_applicable(const rfk_mlp* h);
int fused_prepare(rfk_mlp* h);
void fused_release(rfk_mlp* h);
// out = sum_w probs_w / total_trials over the w trials given (w == total_trials: the mean)
// optional scratch of fused_workspace_bytes() enables the one-item-per-SM schedule for small batches
size_t fused_workspace_bytes(int w, int n_classes, int64_t n, int d);
int fused_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n,
                  float* out, void* ws, size_t ws_bytes, cudaStream_t st);

// device images of a trial packed for the fused kernels (owned by the handle)
struct FusedImages {
  const uint8_t* w1;
  const uint8_t* w2;
  const uint8_t* w3;
  const float* bias;
  int h1_pad, h2_pad, n_hidden;
};
bool fused_images(const rfk_mlp* h, FusedImages* out);

// fused_pair.cu -- CTA-pair kernel for large batches (weights resident in shared memory, trial-outer order)
bool pair_applicable(int d, int64_t n);
size_t pair_workspace_bytes(int w, int n_classes, int64_t n);
int pair_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n, float* out, void* ws,
                 size_t ws_bytes, cudaStream_t st);

// ensemble.cu
int scale_inplace(float* p, int64_t count, double divisor, cudaStream_t st);

}  // namespace rfk
