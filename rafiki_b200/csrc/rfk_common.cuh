// Shared host-side plumbing of librafiki_b200: error reporting and launch helpers.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../../include/rafiki_b200.h"

namespace rfk {

void set_error(const char* fmt, ...);
int num_sms();  // SM count of the current device (cached per device)
void count_launch(const char* kernel_name);  // launch accounting behind rfk_launch_count()

#define RFK_CUDA_TRY(expr)                                                                   \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      rfk::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,       \
                     __LINE__);                                                              \
      return RFK_ERR_CUDA;                                                                   \
    }                                                                                        \
  } while (0)

#define RFK_LAUNCH_CHECK(name)                                                               \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess) {                                                                 \
      rfk::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));               \
      return RFK_ERR_CUDA;                                                                   \
    }                                                                                        \
    rfk::count_launch(name);                                                                 \
  } while (0)

#define RFK_REQUIRE(cond, ...)                                                               \
  do {                                                                                       \
    if (!(cond)) {                                                                           \
      rfk::set_error(__VA_ARGS__);                                                           \
      return RFK_ERR_INVALID;                                                                \
    }                                                                                        \
  } while (0)

// One packed trial (TfFeedForward.py:141-157).  Layer i maps dims[i] -> dims[i+1];
// the last layer is the softmax layer.
struct Layer {
  int in = 0, out = 0;
  // fp32 path: Keras layout [in, out] row-major, bias [out]
  float* w_f32 = nullptr;
  float* b_f32 = nullptr;
  // tensor-core paths: 16-bit weights (bf16 or fp16 bit patterns, per rfk_mlp::dtype),
  // transposed and padded to [out_pad, in_pad] row-major ("K-major"), zero filled; layer 0
  // carries the folded BatchNormalization (see mlp.cu).  Host copies, consumed by the packers.
  int in_pad = 0, out_pad = 0;
  std::vector<uint16_t> w16_host;
  std::vector<float> b_fold_host;  // [out_pad] (layer 0: bias + shift @ W with the ROUNDED weights)
};

}  // namespace rfk

struct rfk_mlp {
  int device = 0;
  int d = 0, n_hidden = 0, n_classes = 0, dtype = RFK_F32;
  int max_units = 0;
  bool fusable = false;
  float* scale = nullptr;  // [d] BN affine (fp32 path and the decode kernel)
  float* shift = nullptr;
  std::vector<rfk::Layer> layers;
  void* fused = nullptr;   // opaque state of the fused tcgen05 path (fused_trial.cu)
  void* tc = nullptr;      // opaque state of the layered tcgen05 path (gemm_tc.cu)
};
