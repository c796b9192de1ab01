// Kernel 3: cross-trial probability averaging.
//
// Replaces the IMAGE_CLASSIFICATION branch of ensemble_predictions
// (rafiki/predictor/ensemble.py:10-14): out[i] = mean_w probs_w[i].  The reference sums
// Python floats (fp32 values widened to float64) in trial order and divides by W in
// float64; this kernel does the same in float64 registers and rounds once to fp32, so
// out == float32(reference) bit for bit.
//
// HBM-bound: (W + 1) * C * 4 algorithmic bytes per query.  16-byte loads from each of the
// W trial buffers (which may be peer-GPU memory), W independent loads in flight per
// thread, persistent grid.
#include <stdlib.h>
#include <string.h>
#include "rfk_internal.cuh"

namespace {

constexpr int kThreads = 256;

struct TrialPtrs {
  const float* p[RFK_MAX_TRIALS];
};

__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// divisor = W for the mean; 1 for a plain float64 sum (partial means of several ranks)
template <int W>
__global__ void __launch_bounds__(kThreads)
ensemble_mean_vec4_kernel(const TrialPtrs ptrs, int w_runtime, int64_t n_vec, float* __restrict__ out, double divisor) {
  // the next kernel of the stream may be a programmatic dependent (the pair kernel of the next batch): let its
  // prologue start while this grid drains; it waits for this grid's completion before it touches global memory
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");       // launched as a programmatic dependent itself (see the launch)
  const int w = W > 0 ? W : w_runtime;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_vec; i += stride) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    if (W > 0) {
      float4 v[W > 0 ? W : 1];
#pragma unroll
      for (int t = 0; t < W; ++t) v[t] = ld_stream_f4(ptrs.p[t] + 4 * i);
#pragma unroll
      for (int t = 0; t < W; ++t) {
        s0 += (double)v[t].x; s1 += (double)v[t].y; s2 += (double)v[t].z; s3 += (double)v[t].w;
      }
    } else {
      for (int t = 0; t < w; ++t) {
        const float4 v = ld_stream_f4(ptrs.p[t] + 4 * i);
        s0 += (double)v.x; s1 += (double)v.y; s2 += (double)v.z; s3 += (double)v.w;
      }
    }
    // np.mean divides the float64 sum by W (true division, not a multiply by 1/W)
    float4 o;
    o.x = (float)(s0 / divisor); o.y = (float)(s1 / divisor);
    o.z = (float)(s2 / divisor); o.w = (float)(s3 / divisor);
    *reinterpret_cast<float4*>(out + 4 * i) = o;
  }
}

__global__ void __launch_bounds__(kThreads)
ensemble_mean_scalar_kernel(const TrialPtrs ptrs, int w, int64_t n_elems, float* __restrict__ out, double divisor) {
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_elems; i += stride) {
    double s = 0.0;
    for (int t = 0; t < w; ++t) s += (double)ptrs.p[t][i];
    out[i] = (float)(s / divisor);
  }
}

}  // namespace

namespace {
int ensemble_reduce(const float* const* probs, int w, int64_t n, int c, float* out, void* stream, double divisor);
}

extern "C" int rfk_ensemble_mean(const float* const* probs, int w, int64_t n, int c, float* out,
                                 void* stream) {
  return ensemble_reduce(probs, w, n, c, out, stream, (double)w);
}

extern "C" int rfk_ensemble_sum(const float* const* probs, int w, int64_t n, int c, float* out, void* stream) {
  return ensemble_reduce(probs, w, n, c, out, stream, 1.0);
}

namespace rfk {
// out = float(sum over the w inputs in float64 / divisor): the partial mean of a rank that holds w of
// `divisor` trials, in one kernel and one rounding
int ensemble_partial_mean(const float* const* probs, int w, int64_t n, int c, float* out, double divisor,
                          cudaStream_t st) {
  return ensemble_reduce(probs, w, n, c, out, st, divisor);
}
}  // namespace rfk

namespace {
int ensemble_reduce(const float* const* probs, int w, int64_t n, int c, float* out, void* stream, double divisor) {
  RFK_REQUIRE(w >= 1 && w <= RFK_MAX_TRIALS, "ensemble_mean: w=%d outside 1..%d", w, RFK_MAX_TRIALS);
  RFK_REQUIRE(n >= 0 && c >= 1, "ensemble_mean: bad shape n=%lld c=%d", (long long)n, c);
  if (n == 0) return RFK_OK;
  RFK_REQUIRE(probs != nullptr && out != nullptr, "ensemble_mean: NULL buffer");
  TrialPtrs ptrs;
  uintptr_t align = reinterpret_cast<uintptr_t>(out);
  for (int t = 0; t < w; ++t) {
    RFK_REQUIRE(probs[t] != nullptr, "ensemble_mean: probs[%d] is NULL", t);
    ptrs.p[t] = probs[t];
    align |= reinterpret_cast<uintptr_t>(probs[t]);
  }
  for (int t = w; t < RFK_MAX_TRIALS; ++t) ptrs.p[t] = nullptr;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t elems = n * (int64_t)c;
  const int sms = rfk::num_sms();
  if (elems % 4 == 0 && (align & 15) == 0) {
    const int64_t n_vec = elems / 4;
    const int64_t want = (n_vec + kThreads - 1) / kThreads;
    const int grid = (int)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    // a programmatic dependent of the kernel before it (normally the trial kernel that wrote `probs`): its
    // blocks are placed as that kernel's CTAs leave and wait, at their top, for the whole grid to complete
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = getenv("RFK_NO_PDL") == nullptr ? 1 : 0;
    cudaError_t le;
    switch (w) {
      case 1: le = cudaLaunchKernelEx(&cfg, ensemble_mean_vec4_kernel<1>, ptrs, w, n_vec, out, divisor); break;
      case 2: le = cudaLaunchKernelEx(&cfg, ensemble_mean_vec4_kernel<2>, ptrs, w, n_vec, out, divisor); break;
      case 3: le = cudaLaunchKernelEx(&cfg, ensemble_mean_vec4_kernel<3>, ptrs, w, n_vec, out, divisor); break;
      case 4: le = cudaLaunchKernelEx(&cfg, ensemble_mean_vec4_kernel<4>, ptrs, w, n_vec, out, divisor); break;
      case 8: le = cudaLaunchKernelEx(&cfg, ensemble_mean_vec4_kernel<8>, ptrs, w, n_vec, out, divisor); break;
      default: le = cudaLaunchKernelEx(&cfg, ensemble_mean_vec4_kernel<0>, ptrs, w, n_vec, out, divisor); break;
    }
    RFK_CUDA_TRY(le);
    RFK_LAUNCH_CHECK("ensemble_mean_vec4_kernel");
  } else {
    const int64_t want = (elems + kThreads - 1) / kThreads;
    const int grid = (int)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    ensemble_mean_scalar_kernel<<<grid, kThreads, 0, st>>>(ptrs, w, elems, out, divisor);
    RFK_LAUNCH_CHECK("ensemble_mean_scalar_kernel");
  }
  return RFK_OK;
}
}  // namespace

namespace {
__global__ void __launch_bounds__(kThreads)
scale_inplace_kernel(float* __restrict__ p, int64_t count, double divisor) {
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < count; i += stride)
    p[i] = (float)((double)p[i] / divisor);
}
}  // namespace

namespace rfk {
int scale_inplace(float* p, int64_t count, double divisor, cudaStream_t st) {
  if (count <= 0) return RFK_OK;
  const int64_t want = (count + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)num_sms() * 8;
  scale_inplace_kernel<<<(int)(want < cap ? want : cap), kThreads, 0, st>>>(p, count, divisor);
  RFK_LAUNCH_CHECK("scale_inplace_kernel");
  return RFK_OK;
}
}  // namespace rfk
