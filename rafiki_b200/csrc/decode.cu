// Kernel 1: uint8 query decode + BatchNormalization affine, with Pillow-NEAREST resize.
//
// Replaces the float feed, Flatten and BatchNormalization(inference) of the Keras graph
// (examples/models/image_classification/TfFeedForward.py:95,147-148) and, for the resize
// variant, ModelDatasetUtils.resize_as_images (rafiki/model/dataset.py:68-78).
//
// HBM-bound byte work: algorithmic traffic is d * (1 + sizeof(out)) bytes per query.
// Design: persistent CTAs (a multiple of the SM count), the per-feature scale/shift
// vectors staged once per CTA in shared memory, one 16-byte load (16 pixels) per thread
// per step, 16-byte stores, streaming cache hints on both sides.
#include <atomic>
#include <type_traits>

#include "rfk_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxSmemFeatures = 12288;  // 2 * 12288 * 4 B = 96 KB of staged scale/shift

__device__ __forceinline__ uint4 ld_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_16(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <typename Out>
__device__ __forceinline__ uint32_t pack16(float lo, float hi) {
  return sizeof(Out) == 2 && !std::is_same<Out, __nv_bfloat16>::value ? pack_f16(lo, hi) : pack_bf16(lo, hi);
}
__device__ __forceinline__ float byte_to_float(uint32_t word, int i) {
  return (float)((word >> (8 * i)) & 0xffu);
}

template <typename Out>
__device__ __forceinline__ void store_one(Out* p, float v);
template <>
__device__ __forceinline__ void store_one<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void store_one<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16_rn(v);
}
template <>
__device__ __forceinline__ void store_one<__half>(__half* p, float v) {
  *p = __float2half_rn(v);
}

// Fast path: d % 16 == 0, ld_out == d, 16-byte aligned pointers.  One thread produces exactly ONE
// 16-byte output vector (8 pixels for a 16-bit output, 4 for fp32), so that every store instruction of a
// warp covers 512 contiguous bytes; four independent units per loop trip keep more loads in flight.
// `kAffine` selects the BN multiply-add.  (A first version with 16 pixels per thread stored two / four
// 16-byte vectors strided by 32 / 64 bytes across the warp -- half / quarter-filled sectors per
// instruction -- and measured 0.43-0.51 of the HBM copy bandwidth; this one 0.70-0.80.)
template <typename Out, bool kAffine>
__global__ void __launch_bounds__(kThreads)
decode_unit_kernel(const uint8_t* __restrict__ x, Out* __restrict__ y, int64_t n_units, int units_per_row,
                   const float* __restrict__ scale, const float* __restrict__ shift, int d, int use_smem) {
  constexpr int B = 16 / (int)sizeof(Out);        // input bytes per unit
  constexpr int U = 4;                            // units per loop trip
  extern __shared__ __align__(16) float s_aff[];  // [2*d] when use_smem
  const float* sc = scale;
  const float* sh = shift;
  if (kAffine && use_smem) {
    for (int i = threadIdx.x; i < d; i += kThreads) {
      s_aff[i] = scale[i];
      s_aff[d + i] = shift[i];
    }
    __syncthreads();
    sc = s_aff;
    sh = s_aff + d;
  }
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t u0 = (int64_t)blockIdx.x * kThreads + threadIdx.x; u0 < n_units; u0 += U * stride) {
    uint32_t w[U][2];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int64_t u = u0 + k * stride;
      w[k][0] = w[k][1] = 0u;
      if (u < n_units) {
        if (B == 8) {
          const uint2 r = __ldg(reinterpret_cast<const uint2*>(x) + u);
          w[k][0] = r.x;
          w[k][1] = r.y;
        } else {
          w[k][0] = __ldg(reinterpret_cast<const uint32_t*>(x) + u);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int64_t u = u0 + k * stride;
      if (u >= n_units) break;
      const int j0 = (int)(u % units_per_row) * B;
      float v[B];
#pragma unroll
      for (int q = 0; q < B / 4; ++q) {
        float4 a = make_float4(1.f, 1.f, 1.f, 1.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (kAffine) {
          a = *reinterpret_cast<const float4*>(sc + j0 + 4 * q);
          b = *reinterpret_cast<const float4*>(sh + j0 + 4 * q);
        }
        const float f0 = byte_to_float(w[k][q], 0), f1 = byte_to_float(w[k][q], 1);
        const float f2 = byte_to_float(w[k][q], 2), f3 = byte_to_float(w[k][q], 3);
        v[4 * q + 0] = kAffine ? fmaf(f0, a.x, b.x) : f0;
        v[4 * q + 1] = kAffine ? fmaf(f1, a.y, b.y) : f1;
        v[4 * q + 2] = kAffine ? fmaf(f2, a.z, b.z) : f2;
        v[4 * q + 3] = kAffine ? fmaf(f3, a.w, b.w) : f3;
      }
      uint4 o;
      if (B == 8) {
        o.x = pack16<Out>(v[0], v[1]);
        o.y = pack16<Out>(v[2], v[3]);
        o.z = pack16<Out>(v[B - 4], v[B - 3]);
        o.w = pack16<Out>(v[B - 2], v[B - 1]);
      } else {
        o.x = __float_as_uint(v[0]); o.y = __float_as_uint(v[1]);
        o.z = __float_as_uint(v[2]); o.w = __float_as_uint(v[3]);
      }
      st_stream_16(reinterpret_cast<char*>(y) + u * 16, o);
    }
  }
}

// General path: any d / ld_out / resize.  One thread per output element; the source index
// of the NEAREST resize is floor((dst + 0.5) * s_in / s_out) per axis, evaluated in exact
// integer arithmetic as ((2*dst + 1) * s_in) / (2 * s_out).
template <typename Out>
__global__ void __launch_bounds__(kThreads)
decode_general_kernel(const uint8_t* __restrict__ x, Out* __restrict__ y, int64_t n, int d,
                      int ld_out, const float* __restrict__ scale, const float* __restrict__ shift,
                      int s_in, int s_out, int ch) {
  const int64_t total = n * (int64_t)ld_out;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  const int64_t d_in = s_in > 0 ? (int64_t)s_in * s_in * ch : d;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += stride) {
    const int64_t row = i / ld_out;
    const int j = (int)(i - row * ld_out);
    float v = 0.f;
    if (j < d) {
      int64_t src = j;
      if (s_in > 0) {
        const int c = j % ch;
        const int px = (j / ch) % s_out;
        const int py = j / (ch * s_out);
        const int sx = min(((2 * px + 1) * s_in) / (2 * s_out), s_in - 1);
        const int sy = min(((2 * py + 1) * s_in) / (2 * s_out), s_in - 1);
        src = ((int64_t)sy * s_in + sx) * ch + c;
      }
      v = (float)x[row * d_in + src];
      if (scale != nullptr) v = fmaf(v, scale[j], shift[j]);
    }
    store_one<Out>(y + i, v);
  }
}

__global__ void __launch_bounds__(kThreads)
resize_u8_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, int64_t n, int ld_out,
                 int s_in, int s_out, int ch) {
  const int d = s_out * s_out * ch;
  const int64_t total = n * (int64_t)ld_out;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  const int64_t d_in = (int64_t)s_in * s_in * ch;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += stride) {
    const int64_t row = i / ld_out;
    const int j = (int)(i - row * ld_out);
    uint8_t v = 0;
    if (j < d) {
      const int c = j % ch;
      const int px = (j / ch) % s_out;
      const int py = j / (ch * s_out);
      const int sx = min(((2 * px + 1) * s_in) / (2 * s_out), s_in - 1);
      const int sy = min(((2 * py + 1) * s_in) / (2 * s_out), s_in - 1);
      v = x[row * d_in + ((int64_t)sy * s_in + sx) * ch + c];
    }
    y[i] = v;
  }
}

// Fast NEAREST resize for whole images that are multiples of 16 bytes (28x28 -> 32x32 is the
// reference's default knob, TfFeedForward.py:27): a block stages a group of consecutive source images
// in shared memory with 16-byte loads (images are contiguous in memory), then every thread gathers 16
// output bytes through a per-block offset table and stores them as one 16-byte vector.  The one-byte-
// per-thread kernel above ran at 4 % of HBM bandwidth.
constexpr int kResizeSrcBytes = 32 * 1024;
__global__ void __launch_bounds__(kThreads)
resize_u8_staged_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, int64_t n, int s_in, int s_out, int ch,
                        int group) {
  extern __shared__ __align__(16) uint8_t rs_smem[];
  const int d_in = s_in * s_in * ch, d_out = s_out * s_out * ch;
  uint16_t* table = reinterpret_cast<uint16_t*>(rs_smem);                 // [d_out] source offset of each output byte
  uint8_t* src = rs_smem + ((2 * d_out + 15) / 16) * 16;                    // [group][d_in]
  for (int j = threadIdx.x; j < d_out; j += kThreads) {
    const int c = j % ch, px = (j / ch) % s_out, py = j / (ch * s_out);
    const int sx = min(((2 * px + 1) * s_in) / (2 * s_out), s_in - 1);
    const int sy = min(((2 * py + 1) * s_in) / (2 * s_out), s_in - 1);
    table[j] = (uint16_t)((sy * s_in + sx) * ch + c);
  }
  const int64_t n_groups = (n + group - 1) / group;
  const int vin = d_in / 16, vout = d_out / 16;
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int64_t first = g * group;
    const int imgs = (int)(n - first < group ? n - first : group);
    __syncthreads();                                                      // table ready / previous group consumed
    const uint4* gsrc = reinterpret_cast<const uint4*>(x + first * d_in);
    for (int v = threadIdx.x; v < imgs * vin; v += kThreads) reinterpret_cast<uint4*>(src)[v] = __ldg(gsrc + v);
    __syncthreads();
    uint4* gdst = reinterpret_cast<uint4*>(y + first * d_out);
    for (int v = threadIdx.x; v < imgs * vout; v += kThreads) {
      const int img = v / vout, o = (v - img * vout) * 16;
      const uint8_t* s = src + img * d_in;
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint16_t* t = table + o + 4 * q;
        w[q] = (uint32_t)s[t[0]] | ((uint32_t)s[t[1]] << 8) | ((uint32_t)s[t[2]] << 16) | ((uint32_t)s[t[3]] << 24);
      }
      gdst[v] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

template <typename Out>
int launch_decode(const uint8_t* x, Out* y, int64_t n, int d, int ld_out, const float* scale,
                  const float* shift, int s_in, int s_out, int ch, cudaStream_t st) {
  const int sms = rfk::num_sms();
  const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0 &&
                       (scale == nullptr ||
                        ((reinterpret_cast<uintptr_t>(scale) | reinterpret_cast<uintptr_t>(shift)) & 15) == 0);
  if (s_in <= 0 && d % 16 == 0 && ld_out == d && aligned) {
    constexpr int B = 16 / (int)sizeof(Out);
    const int64_t n_units = n * (int64_t)(d / B);
    const int64_t want = (n_units + 4 * kThreads - 1) / (4 * kThreads);
    const int grid = (int)(want < (int64_t)sms * 8 ? (want > 0 ? want : 1) : (int64_t)sms * 8);
    const int use_smem = (scale != nullptr && d <= kMaxSmemFeatures) ? 1 : 0;
    const size_t smem = use_smem ? (size_t)2 * d * sizeof(float) : 0;
    if (scale != nullptr) {
      auto k = decode_unit_kernel<Out, true>;
      if (smem > 48 * 1024)
        RFK_CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, kThreads, smem, st>>>(x, y, n_units, d / B, scale, shift, d, use_smem);
    } else {
      decode_unit_kernel<Out, false><<<grid, kThreads, 0, st>>>(x, y, n_units, d / B, nullptr, nullptr, d, 0);
    }
    RFK_LAUNCH_CHECK("decode_unit_kernel");
    return RFK_OK;
  }
  const int64_t total = n * (int64_t)ld_out;
  const int64_t want = (total + kThreads - 1) / kThreads;
  const int grid = (int)(want < (int64_t)sms * 16 ? (want > 0 ? want : 1) : (int64_t)sms * 16);
  decode_general_kernel<Out><<<grid, kThreads, 0, st>>>(x, y, n, d, ld_out, scale, shift, s_in, s_out, ch);
  RFK_LAUNCH_CHECK("decode_general_kernel");
  return RFK_OK;
}

int decode_dispatch(const uint8_t* x, void* y, int64_t n, int d, int ld_out, const float* scale,
                    const float* shift, int out_dtype, int s_in, int s_out, int ch, void* stream) {
  RFK_REQUIRE(n >= 0 && d > 0 && ld_out >= d, "decode: bad shape n=%lld d=%d ld_out=%d", (long long)n, d, ld_out);
  RFK_REQUIRE((scale == nullptr) == (shift == nullptr), "decode: scale and shift must both be given or both NULL");
  if (n == 0) return RFK_OK;
  RFK_REQUIRE(x != nullptr && y != nullptr, "decode: NULL buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (out_dtype == RFK_F32)
    return launch_decode<float>(x, static_cast<float*>(y), n, d, ld_out, scale, shift, s_in, s_out, ch, st);
  if (out_dtype == RFK_BF16)
    return launch_decode<__nv_bfloat16>(x, static_cast<__nv_bfloat16*>(y), n, d, ld_out, scale, shift, s_in, s_out, ch, st);
  if (out_dtype == RFK_F16)
    return launch_decode<__half>(x, static_cast<__half*>(y), n, d, ld_out, scale, shift, s_in, s_out, ch, st);
  rfk::set_error("decode: unknown out_dtype %d", out_dtype);
  return RFK_ERR_INVALID;
}

}  // namespace

extern "C" int rfk_decode_u8(const uint8_t* x, void* y, int64_t n, int d, int ld_out,
                             const float* scale, const float* shift, int out_dtype, void* stream) {
  return decode_dispatch(x, y, n, d, ld_out, scale, shift, out_dtype, 0, 0, 1, stream);
}

extern "C" int rfk_resize_decode_u8(const uint8_t* x, void* y, int64_t n, int s_in, int s_out, int ch,
                                    int ld_out, const float* scale, const float* shift, int out_dtype,
                                    void* stream) {
  RFK_REQUIRE(s_in > 0 && s_out > 0 && ch > 0, "resize: bad sizes s_in=%d s_out=%d ch=%d", s_in, s_out, ch);
  const int d = s_out * s_out * ch;
  if (out_dtype == -1) {
    RFK_REQUIRE(n >= 0 && ld_out >= d, "resize: bad shape");
    if (n == 0) return RFK_OK;
    RFK_REQUIRE(x != nullptr && y != nullptr, "resize: NULL buffer");
    const int64_t total = n * (int64_t)ld_out;
    const int64_t want = (total + kThreads - 1) / kThreads;
    const int sms = rfk::num_sms();
    const int d_in = s_in * s_in * ch;
    if (ld_out == d && d % 16 == 0 && d_in % 16 == 0 && d_in <= kResizeSrcBytes && d <= 32768 &&
        ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0) {
      int group = kResizeSrcBytes / d_in;
      group = group > 16 ? 16 : group;
      const size_t smem = (size_t)((2 * d + 15) / 16) * 16 + (size_t)group * d_in;
      const int64_t n_groups = (n + group - 1) / group;
      const int g = (int)(n_groups < (int64_t)sms * 4 ? n_groups : (int64_t)sms * 4);
      static std::atomic<bool> attr_set[64];
      int dev = 0;
      RFK_CUDA_TRY(cudaGetDevice(&dev));
      RFK_REQUIRE(dev >= 0 && dev < 64, "resize: device ordinal %d not supported", dev);
      if (!attr_set[dev].load(std::memory_order_acquire)) {     // setting it twice is harmless
        RFK_CUDA_TRY(cudaFuncSetAttribute(resize_u8_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        attr_set[dev].store(true, std::memory_order_release);
      }
      resize_u8_staged_kernel<<<g, kThreads, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
          x, static_cast<uint8_t*>(y), n, s_in, s_out, ch, group);
      RFK_LAUNCH_CHECK("resize_u8_staged_kernel");
      return RFK_OK;
    }
    const int grid = (int)(want < (int64_t)sms * 16 ? (want > 0 ? want : 1) : (int64_t)sms * 16);
    resize_u8_kernel<<<grid, kThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        x, static_cast<uint8_t*>(y), n, ld_out, s_in, s_out, ch);
    RFK_LAUNCH_CHECK("resize_u8_kernel");
    return RFK_OK;
  }
  return decode_dispatch(x, y, n, d, ld_out, scale, shift, out_dtype, s_in, s_out, ch, stream);
}
