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
  constexpr int U = B == 8 ? 8 : 4;               // units per loop trip (loads in flight per thread)
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
// When the image is enlarged, four consecutive output bytes come from at most eight consecutive source
// bytes, so an output word is ONE byte permute of two aligned source words: the block keeps a second table
// (source word index | PRMT selector << 16 per output word) and a thread builds its 16 output bytes with
// one 16-byte table load, eight 4-byte loads and four permutes instead of sixteen byte gathers (the byte
// gathers made the kernel shared-memory-issue bound at about half of the HBM rate).  A block whose table
// has an output word that does not fit (a reduction, many channels) falls back to the byte gathers.
__global__ void __launch_bounds__(kThreads)
resize_u8_staged_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, int64_t n, int s_in, int s_out, int ch,
                        int group) {
  extern __shared__ __align__(16) uint8_t rs_smem[];
  const int d_in = s_in * s_in * ch, d_out = s_out * s_out * ch;
  uint16_t* table = reinterpret_cast<uint16_t*>(rs_smem);                 // [d_out] source offset of each output byte
  uint32_t* wtab = reinterpret_cast<uint32_t*>(rs_smem + ((2 * d_out + 15) / 16) * 16);   // [d_out / 4]
  uint8_t* src = reinterpret_cast<uint8_t*>(wtab) + d_out;                // 2 x ([group][d_in] + 16 bytes of slack)
  for (int j = threadIdx.x; j < d_out; j += kThreads) {
    const int c = j % ch, px = (j / ch) % s_out, py = j / (ch * s_out);
    const int sx = min(((2 * px + 1) * s_in) / (2 * s_out), s_in - 1);
    const int sy = min(((2 * py + 1) * s_in) / (2 * s_out), s_in - 1);
    table[j] = (uint16_t)((sy * s_in + sx) * ch + c);
  }
  __syncthreads();
  int bad = 0;
  for (int w = threadIdx.x; w < d_out / 4; w += kThreads) {
    const int o0 = table[4 * w], o1 = table[4 * w + 1], o2 = table[4 * w + 2], o3 = table[4 * w + 3];
    const int base = min(min(o0, o1), min(o2, o3)) & ~3;
    const int hi = max(max(o0, o1), max(o2, o3));
    if (hi - base >= 8) bad = 1;
    wtab[w] = (uint32_t)(base >> 2) |
              ((uint32_t)((o0 - base) | ((o1 - base) << 4) | ((o2 - base) << 8) | ((o3 - base) << 12)) << 16);
  }
  const bool by_words = __syncthreads_or(bad) == 0;
  const int64_t n_groups = (n + group - 1) / group;
  const int vin = d_in / 16, vout = d_out / 16;
  // Two source buffers: the 16-byte asynchronous copies of the NEXT group are in flight while this one is
  // gathered, so a block never sits between "loaded" and "stored" with nothing outstanding.
  const uint32_t buf_bytes = (uint32_t)group * (uint32_t)d_in + 16u;
  auto prefetch = [&](int64_t g, int b) {
    const int64_t first = g * group;
    const int imgs = (int)(n - first < group ? n - first : group);
    const uint8_t* gsrc = x + first * d_in;
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(src + (size_t)b * buf_bytes);
    for (int v = threadIdx.x; v < imgs * vin; v += kThreads)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * v), "l"(gsrc + 16 * (size_t)v) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int cur = 0;
  if ((int64_t)blockIdx.x < n_groups) prefetch(blockIdx.x, 0);
  for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x, cur ^= 1) {
    const int64_t first = g * group;
    const int imgs = (int)(n - first < group ? n - first : group);
    if (g + gridDim.x < n_groups) {
      prefetch(g + gridDim.x, cur ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();                                                      // this group's bytes (and the tables) are visible
    const uint8_t* sbuf = src + (size_t)cur * buf_bytes;
    uint4* gdst = reinterpret_cast<uint4*>(y + first * d_out);
    if (by_words) {
      for (int v = threadIdx.x; v < imgs * vout; v += kThreads) {
        const int img = v / vout, ov = v - img * vout;
        const uint32_t* s32 = reinterpret_cast<const uint32_t*>(sbuf + img * d_in);
        const uint4 e = reinterpret_cast<const uint4*>(wtab)[ov];
        uint4 o;
        o.x = __byte_perm(s32[e.x & 0xffffu], s32[(e.x & 0xffffu) + 1], e.x >> 16);
        o.y = __byte_perm(s32[e.y & 0xffffu], s32[(e.y & 0xffffu) + 1], e.y >> 16);
        o.z = __byte_perm(s32[e.z & 0xffffu], s32[(e.z & 0xffffu) + 1], e.z >> 16);
        o.w = __byte_perm(s32[e.w & 0xffffu], s32[(e.w & 0xffffu) + 1], e.w >> 16);
        st_stream_16(reinterpret_cast<char*>(gdst + v), o);
      }
    } else {
      for (int v = threadIdx.x; v < imgs * vout; v += kThreads) {
        const int img = v / vout, o = (v - img * vout) * 16;
        const uint8_t* s = sbuf + img * d_in;
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint16_t* t = table + o + 4 * q;
          w[q] = (uint32_t)s[t[0]] | ((uint32_t)s[t[1]] << 8) | ((uint32_t)s[t[2]] << 16) | ((uint32_t)s[t[3]] << 24);
        }
        gdst[v] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
    __syncthreads();                                                      // done with this buffer before it is refilled
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
    const int64_t want = (n_units + (B == 8 ? 8 : 4) * kThreads - 1) / ((B == 8 ? 8 : 4) * kThreads);
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
    if (ld_out == d && d % 16 == 0 && d_in % 16 == 0 && 2 * d_in <= kResizeSrcBytes && d <= 32768 &&
        ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0) {
      int group = kResizeSrcBytes / (2 * d_in);                 // two buffers
      group = group > 8 ? 8 : (group < 1 ? 1 : group);
      const size_t smem = (size_t)((2 * d + 15) / 16) * 16 + (size_t)d + 2 * ((size_t)group * d_in + 16);
      const int64_t n_groups = (n + group - 1) / group;
      const int g = (int)(n_groups < (int64_t)sms * 8 ? n_groups : (int64_t)sms * 8);
      static std::atomic<bool> attr_set[64];
      int dev = 0;
      RFK_CUDA_TRY(cudaGetDevice(&dev));
      RFK_REQUIRE(dev >= 0 && dev < 64, "resize: device ordinal %d not supported", dev);
      if (!attr_set[dev].load(std::memory_order_acquire)) {     // setting it twice is harmless
        RFK_CUDA_TRY(cudaFuncSetAttribute(resize_u8_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 140 * 1024));
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
