// Kernel 2 host side: packing a trial (load_parameters) and dispatching its forward pass.
//
// Replaces TfFeedForward.load_parameters / predict (examples/models/image_classification/
// TfFeedForward.py:89-97,118-135): rfk_mlp_create takes the Keras tensors as host arrays
// (Dense kernels [in, out], biases, BatchNormalization gamma/beta/moving mean/variance) and
// keeps device copies in the layouts the kernels want.
//
// BatchNormalization folding (tensor-core paths, 16-bit weights w' in bf16 or fp16).  With inv = gamma * rsqrt(var + eps),
//   y_j = sum_i ((x_i - mean_i) * inv_i + beta_i) * K_ij + b_j
//       = sum_i x_i * (inv_i * K_ij) + [b_j + sum_i beta_i * K_ij - sum_i mean_i * inv_i * K_ij].
// The uint8 pixels x_i are exact in bf16 and in fp16, so the only rounding is w'_ij = round16(inv_i * K_ij).
// The constant uses the ROUNDED weights, c_j = b_j + sum_i beta_i K_ij - sum_i mean_i w'_ij,
// which makes the error sum_i (x_i - mean_i) * (w'_ij - inv_i K_ij): it scales with the
// centred pixel, exactly as rounding the normalised activation would, but with one rounding
// instead of two.
#include <math.h>

#include "rfk_internal.cuh"

namespace {

template <typename T>
int upload(T** dst, const T* src, size_t count) {
  RFK_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(dst), count * sizeof(T)));
  RFK_CUDA_TRY(cudaMemcpy(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice));
  return RFK_OK;
}

int round_up(int v, int m) { return (v + m - 1) / m * m; }

void free_all(rfk_mlp* h) {
  if (h == nullptr) return;
  rfk::fused_release(h);
  rfk::tc_release(h);
  cudaFree(h->scale);
  cudaFree(h->shift);
  for (rfk::Layer& l : h->layers) {
    cudaFree(l.w_f32);
    cudaFree(l.b_f32);
  }
  delete h;
}

}  // namespace

extern "C" int rfk_mlp_create(rfk_mlp_t** out, int d, int n_hidden, const int* units, int n_classes,
                              const float* const* kernels, const float* const* biases,
                              const float* bn_gamma, const float* bn_beta, const float* bn_mean,
                              const float* bn_var, float bn_eps, int compute_dtype) {
  RFK_REQUIRE(out != nullptr, "mlp_create: out is NULL");
  *out = nullptr;
  RFK_REQUIRE(d >= 1 && n_hidden >= 0 && n_hidden <= 16 && n_classes >= 1,
              "mlp_create: bad shape d=%d n_hidden=%d n_classes=%d", d, n_hidden, n_classes);
  RFK_REQUIRE(n_hidden == 0 || units != nullptr, "mlp_create: units is NULL");
  RFK_REQUIRE(kernels != nullptr && biases != nullptr, "mlp_create: kernels/biases is NULL");
  RFK_REQUIRE(bn_gamma && bn_beta && bn_mean && bn_var, "mlp_create: a BatchNormalization vector is NULL");
  RFK_REQUIRE(compute_dtype == RFK_F32 || compute_dtype == RFK_BF16 || compute_dtype == RFK_F16,
              "mlp_create: unknown compute_dtype %d", compute_dtype);
  for (int i = 0; i < n_hidden; ++i) RFK_REQUIRE(units[i] >= 1, "mlp_create: units[%d]=%d", i, units[i]);
  for (int i = 0; i <= n_hidden; ++i)
    RFK_REQUIRE(kernels[i] != nullptr && biases[i] != nullptr, "mlp_create: layer %d tensors are NULL", i);
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  int rc = rfk_device_check(dev);
  if (rc != RFK_OK) return rc;

  rfk_mlp* h = new rfk_mlp();
  h->device = dev;
  h->d = d;
  h->n_hidden = n_hidden;
  h->n_classes = n_classes;
  h->dtype = compute_dtype;

  // BatchNormalization affine, evaluated in fp32 exactly like the oracle's bn_affine():
  // inv = gamma / sqrt(var + eps); shift = beta - mean * inv
  std::vector<float> scale(d), shift(d);
  for (int i = 0; i < d; ++i) {
    volatile float ve = bn_var[i] + bn_eps;
    volatile float sq = sqrtf(ve);
    volatile float inv = bn_gamma[i] / sq;
    volatile float mi = bn_mean[i] * inv;
    scale[i] = inv;
    shift[i] = bn_beta[i] - mi;
  }
  rc = upload(&h->scale, scale.data(), d);
  if (rc == RFK_OK) rc = upload(&h->shift, shift.data(), d);

  h->layers.resize(n_hidden + 1);
  int in = d;
  for (int i = 0; i <= n_hidden && rc == RFK_OK; ++i) {
    rfk::Layer& l = h->layers[i];
    l.in = in;
    l.out = i < n_hidden ? units[i] : n_classes;
    if (i < n_hidden && l.out > h->max_units) h->max_units = l.out;
    l.in_pad = round_up(l.in, 64);
    l.out_pad = round_up(l.out, 16);
    if (compute_dtype == RFK_F32) {
      rc = upload(&l.w_f32, kernels[i], (size_t)l.in * l.out);
      if (rc == RFK_OK) rc = upload(&l.b_f32, biases[i], (size_t)l.out);
    } else {
      std::vector<uint16_t> wt((size_t)l.out_pad * l.in_pad, 0);
      std::vector<float> bias(l.out_pad, 0.f);
      for (int j = 0; j < l.out && rc == RFK_OK; ++j) {
        double c = (double)biases[i][j];
        for (int k = 0; k < l.in; ++k) {
          const double kij = (double)kernels[i][(size_t)k * l.out + j];
          double w = kij;
          if (i == 0) {
            const double inv = (double)bn_gamma[k] / sqrt((double)bn_var[k] + (double)bn_eps);
            w = inv * kij;
          }
          double wr;  // the weight as the tensor core will see it
          if (compute_dtype == RFK_F16) {
            if (!(fabs(w) <= 65504.0)) {
              rfk::set_error("mlp_create: layer %d weight [%d,%d] = %g does not fit fp16; use RFK_BF16 or RFK_F32",
                             i, k, j, w);
              rc = RFK_ERR_UNSUPPORTED;
              break;
            }
            const __half hv = __float2half_rn((float)w);
            wt[(size_t)j * l.in_pad + k] = __half_as_ushort(hv);
            wr = (double)__half2float(hv);
          } else {
            const __nv_bfloat16 bv = __float2bfloat16_rn((float)w);
            wt[(size_t)j * l.in_pad + k] = __bfloat16_as_ushort(bv);
            wr = (double)__bfloat162float(bv);
          }
          if (i == 0) c += (double)bn_beta[k] * kij - (double)bn_mean[k] * wr;
        }
        bias[j] = (float)c;
      }
      l.w16_host.swap(wt);
      l.b_fold_host.swap(bias);
    }
    in = l.out;
  }
  if (rc == RFK_OK && compute_dtype != RFK_F32) {
    h->fusable = rfk::fused_applicable(h);
    rc = h->fusable ? rfk::fused_prepare(h) : rfk::tc_prepare(h);
  }
  if (rc != RFK_OK) {
    free_all(h);
    return rc;
  }
  *out = h;
  return RFK_OK;
}

extern "C" int rfk_mlp_destroy(rfk_mlp_t* h) {
  if (h == nullptr) return RFK_OK;
  int cur = 0;
  cudaGetDevice(&cur);
  if (cur != h->device) cudaSetDevice(h->device);
  free_all(h);
  if (cur != h->device) cudaSetDevice(cur);
  return RFK_OK;
}

extern "C" int rfk_mlp_info(const rfk_mlp_t* h, int* out6) {
  RFK_REQUIRE(h != nullptr && out6 != nullptr, "mlp_info: NULL argument");
  out6[0] = h->d;
  out6[1] = h->n_hidden;
  out6[2] = h->n_classes;
  out6[3] = h->dtype;
  out6[4] = h->fusable ? 1 : 0;
  out6[5] = h->max_units;
  return RFK_OK;
}

extern "C" int rfk_mlp_workspace_bytes(const rfk_mlp_t* h, int64_t n, size_t* out) {
  RFK_REQUIRE(h != nullptr && out != nullptr && n >= 0, "mlp_workspace_bytes: bad argument");
  if (h->dtype == RFK_F32)
    *out = rfk::mlp_f32_workspace_bytes(h, n);
  else
    *out = h->fusable ? 0 : rfk::tc_workspace_bytes(h, n);
  return RFK_OK;
}

extern "C" int rfk_mlp_forward(rfk_mlp_t* h, const uint8_t* x, int64_t n, float* probs, void* workspace,
                               size_t workspace_bytes, void* stream) {
  RFK_REQUIRE(h != nullptr, "mlp_forward: handle is NULL");
  RFK_REQUIRE(n >= 0, "mlp_forward: n=%lld", (long long)n);
  if (n == 0) return RFK_OK;
  RFK_REQUIRE(x != nullptr && probs != nullptr, "mlp_forward: NULL buffer");
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  RFK_REQUIRE(dev == h->device, "mlp_forward: trial lives on device %d, current device is %d", h->device, dev);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (h->dtype == RFK_F32) return rfk::mlp_f32_forward(h, x, n, probs, workspace, workspace_bytes, st);
  if (h->fusable) {
    rfk_mlp* one[1] = {h};
    return rfk::fused_forward(one, 1, 1, x, n, probs, nullptr, 0, st);
  }
  return rfk::tc_forward(h, x, n, probs, workspace, workspace_bytes, st);
}

// ------------------------------------------------------------------------------- ensemble
namespace {

bool all_fusable(rfk_mlp_t* const* trials, int w) {
  for (int t = 0; t < w; ++t)
    if (!(trials[t]->dtype == trials[0]->dtype && trials[t]->dtype != RFK_F32 && trials[t]->fusable)) return false;
  return true;
}

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

extern "C" int rfk_ensemble_workspace_bytes(rfk_mlp_t* const* trials, int w, int64_t n, size_t* out) {
  RFK_REQUIRE(trials != nullptr && out != nullptr && w >= 1 && w <= RFK_MAX_TRIALS && n >= 0,
              "ensemble_workspace_bytes: bad argument");
  for (int t = 0; t < w; ++t) RFK_REQUIRE(trials[t] != nullptr, "ensemble_workspace_bytes: trials[%d] is NULL", t);
  if (all_fusable(trials, w)) {
    *out = rfk::fused_workspace_bytes(w, trials[0]->n_classes, n, trials[0]->d);
    return RFK_OK;
  }
  // unfused: one probs buffer per trial + the largest per-trial scratch (trials run in turn)
  size_t scratch = 0;
  for (int t = 0; t < w; ++t) {
    size_t s = 0;
    int rc = rfk_mlp_workspace_bytes(trials[t], n, &s);
    if (rc != RFK_OK) return rc;
    scratch = s > scratch ? s : scratch;
  }
  *out = (size_t)w * align256((size_t)n * trials[0]->n_classes * sizeof(float)) + align256(scratch);
  return RFK_OK;
}

extern "C" int rfk_ensemble_forward(rfk_mlp_t* const* trials, int w, int total_trials, const uint8_t* x,
                                    int64_t n, float* out, void* workspace, size_t workspace_bytes,
                                    void* stream) {
  RFK_REQUIRE(trials != nullptr && w >= 1 && w <= RFK_MAX_TRIALS, "ensemble_forward: w=%d outside 1..%d", w,
              RFK_MAX_TRIALS);
  RFK_REQUIRE(total_trials >= w, "ensemble_forward: total_trials=%d < w=%d", total_trials, w);
  RFK_REQUIRE(n >= 0, "ensemble_forward: n=%lld", (long long)n);
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  for (int t = 0; t < w; ++t) {
    RFK_REQUIRE(trials[t] != nullptr, "ensemble_forward: trials[%d] is NULL", t);
    RFK_REQUIRE(trials[t]->d == trials[0]->d && trials[t]->n_classes == trials[0]->n_classes,
                "ensemble_forward: trial %d has a different input size or class count", t);
    RFK_REQUIRE(trials[t]->device == dev, "ensemble_forward: trial %d lives on device %d, current is %d", t,
                trials[t]->device, dev);
  }
  if (n == 0) return RFK_OK;
  RFK_REQUIRE(x != nullptr && out != nullptr, "ensemble_forward: NULL buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (all_fusable(trials, w))
    return rfk::fused_forward(trials, w, total_trials, x, n, out, workspace, workspace_bytes, st);

  size_t need = 0;
  int rc = rfk_ensemble_workspace_bytes(trials, w, n, &need);
  if (rc != RFK_OK) return rc;
  RFK_REQUIRE(workspace != nullptr && workspace_bytes >= need, "ensemble_forward: workspace too small (%zu < %zu)",
              workspace_bytes, need);
  const int c = trials[0]->n_classes;
  const size_t one = align256((size_t)n * c * sizeof(float));
  char* base = static_cast<char*>(workspace);
  char* scratch = base + (size_t)w * one;
  const size_t scratch_bytes = workspace_bytes - (size_t)w * one;
  const float* ptrs[RFK_MAX_TRIALS];
  for (int t = 0; t < w; ++t) {
    float* p = reinterpret_cast<float*>(base + (size_t)t * one);
    rc = rfk_mlp_forward(trials[t], x, n, p, scratch, scratch_bytes, stream);
    if (rc != RFK_OK) return rc;
    ptrs[t] = p;
  }
  rc = rfk_ensemble_mean(ptrs, w, n, c, out, stream);
  if (rc != RFK_OK) return rc;
  if (total_trials != w) return rfk::scale_inplace(out, n * (int64_t)c, (double)total_trials / (double)w, st);
  return RFK_OK;
}
