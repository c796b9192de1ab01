// Device helpers shared by the fused trial kernels (fused_trial.cu: one CTA per SM, weights streamed;
// fused_pair.cu: CTA pairs, weights resident): exact uint8 -> 16-bit operand conversions, the
// bias + ReLU + pack epilogue that writes hidden activations back into tensor memory, the softmax
// of the last layer, barrier helpers.
#pragma once
#include "sm100.cuh"

namespace fusedk {

using namespace sm100;

// exact uint8 -> bf16: float(b) = as_float(0x4B000000 | b) - 2^23 has at most 8 significant
// bits, so its upper 16 bits ARE the bf16 value (no rounding).
__device__ __forceinline__ uint32_t u8x2_to_bf16x2(uint32_t word, int e0) {
  const float f0 = __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | e0)) - 8388608.0f;
  const float f1 = __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | (e0 + 1))) - 8388608.0f;
  return __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632u);
}
// exact uint8 -> fp16: 0x6400 | b is the half 1024 + b (ulp 1 in [1024, 2048)); subtract 1024.
__device__ __forceinline__ uint32_t u8x2_to_f16x2(uint32_t word, int e0) {
  const uint32_t biased = __byte_perm(word, 0x64646464u, 0x4040u | (uint32_t)e0 | ((uint32_t)(e0 + 1) << 8));
  uint32_t r;
  asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(biased), "r"(0x64006400u));
  return r;
}
// uint8 -> fp16 SUBNORMAL: the 16-bit pattern 0x00bb is the half-precision value b * 2^-24 (no implicit
// bit), so zero-extending the bytes IS the conversion -- one PRMT per pair and no arithmetic.  The
// tensor core multiplies subnormal operands exactly; the layer-1 epilogue scales the fp32
// accumulator back by 2^24 inside the FFMA that adds the bias (a power of two: bit-identical to
// converting to b itself).
__device__ __forceinline__ uint32_t u8x2_to_f16x2_subnormal(uint32_t word, int e0) {
  return __byte_perm(word, 0u, 0x4040u | (uint32_t)e0 | ((uint32_t)(e0 + 1) << 8));
}
constexpr float kSubnormalScale = 16777216.0f;   // 2^24
template <bool kF16>
__device__ __forceinline__ uint32_t u8x2_to_op(uint32_t word, int e0) {
  return kF16 ? u8x2_to_f16x2_subnormal(word, e0) : u8x2_to_bf16x2(word, e0);
}

// relu(lo), relu(hi) -> packed 16-bit operand pair (lo in the low half), one instruction;
// fp16 saturates at 65504 instead of overflowing to infinity
template <bool kF16>
__device__ __forceinline__ uint32_t relu_pack_x2(float lo, float hi) {
  uint32_t r;
  if (kF16)
    asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  else
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

__device__ __forceinline__ uint4 lds_u4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts_u4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// 16 accumulator columns of one row -> bias + ReLU -> 16-bit pairs -> 8 columns of the hidden
// activations in tensor memory (element k of the row lives in the low/high half of column k/2)
template <bool kF16>
__device__ __forceinline__ void hidden_store16(const uint32_t (&v)[16], int c, const float* bias, float scale,
                                               uint32_t hid_row_addr) {
  const float4* b4 = reinterpret_cast<const float4*>(bias + c);
  uint32_t p[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float4 bb = b4[j];
    p[2 * j] = relu_pack_x2<kF16>(fmaf(__uint_as_float(v[4 * j + 0]), scale, bb.x),
                                  fmaf(__uint_as_float(v[4 * j + 1]), scale, bb.y));
    p[2 * j + 1] = relu_pack_x2<kF16>(fmaf(__uint_as_float(v[4 * j + 2]), scale, bb.z),
                                      fmaf(__uint_as_float(v[4 * j + 3]), scale, bb.w));
  }
  tmem_st8(hid_row_addr + (uint32_t)(c >> 1), p);
}

// Columns [c_begin, c_end) of one TMEM accumulator row -> hidden activations (A operand of the
// next layer, kept in tensor memory).  16 columns per step; the TMEM load of step s+1 is in flight
// while step s is converted and stored (two statically named buffers: everything stays in registers;
// four buffers were measured to spill under the 96-register launch bound).
template <bool kF16>
__device__ __forceinline__ void epilogue_hidden(uint32_t tmem_row_addr, int c_begin, int c_end, const float* bias,
                                                float scale, uint32_t hid_row_addr) {
  if (c_begin < c_end) {
    uint32_t va[16], vb[16];
    tmem_ld16(tmem_row_addr + (uint32_t)c_begin, va);
    for (int c = c_begin; c < c_end; c += 32) {
      tmem_ld_wait();
      const bool more_b = c + 16 < c_end;
      if (more_b) tmem_ld16(tmem_row_addr + (uint32_t)c + 16u, vb);
      hidden_store16<kF16>(va, c, bias, scale, hid_row_addr);
      if (more_b) {
        tmem_ld_wait();
        if (c + 32 < c_end) tmem_ld16(tmem_row_addr + (uint32_t)c + 32u, va);
        hidden_store16<kF16>(vb, c + 16, bias, scale, hid_row_addr);
      }
    }
  }
  tmem_st_wait();
}

// Hardware named barriers (bar.sync / bar.arrive): producers arrive without blocking, the consumer
// syncs; `threads` counts both sides and is a multiple of 32.  Every lane of a warp executes them.
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// One arrival per warp instead of one per thread: 32 arrivals on one mbarrier serialise in the
// shared-memory atomic unit (~1 lane per cycle).  __syncwarp orders the lanes' earlier writes
// (and their proxy / tcgen05 fences) before the elected lane's release-arrive.
__device__ __forceinline__ void warp_arrive(uint64_t* bar) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
}
__device__ __forceinline__ void warp_arrive2(uint64_t* bar_a, uint64_t* bar_b) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) {
    mbar_arrive(bar_a);
    mbar_arrive(bar_b);
  }
}

// Last layer: 16 accumulator columns + bias -> softmax.  z[j] = exp(logit_j - max), returns 1 / sum.
// The padded classes need no predicate: their bias entry in the packed image is -infinity (fused_prepare)
// and their weight rows are zero, so their logit is -infinity and their exponential an exact zero.  The
// whole function is a serial chain on one warp per SM sub-partition (about 5 cycles per instruction), so it
// is kept short: one FFMA + ex2.approx per class, one approximate reciprocal; both are far inside the
// tolerance of the 16-bit path (2e-3 on a probability) and identical in the two fused kernels.
__device__ __forceinline__ float softmax16(const uint32_t (&v3)[16], const float* b3, int /*C*/, float (&z)[16]) {
  const float4* b4 = reinterpret_cast<const float4*>(b3);
  float mx = -INFINITY;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 bb = b4[q];
    z[4 * q + 0] = __uint_as_float(v3[4 * q + 0]) + bb.x;
    z[4 * q + 1] = __uint_as_float(v3[4 * q + 1]) + bb.y;
    z[4 * q + 2] = __uint_as_float(v3[4 * q + 2]) + bb.z;
    z[4 * q + 3] = __uint_as_float(v3[4 * q + 3]) + bb.w;
  }
#pragma unroll
  for (int j = 0; j < 16; j += 2) mx = fmaxf(mx, fmaxf(z[j], z[j + 1]));
  const float kLog2e = 1.4426950408889634f;
  const float off = -mx * kLog2e;
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int j = 0; j < 16; j += 2) {
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(z[j]) : "f"(fmaf(z[j], kLog2e, off)));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(z[j + 1]) : "f"(fmaf(z[j + 1], kLog2e, off)));
    s0 += z[j];
    s1 += z[j + 1];
  }
  float inv;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(s0 + s1));
  return inv;
}
// predicated 4-byte shared-memory store (no branch)
__device__ __forceinline__ void sts_f32_if(uint32_t addr, float v, bool pred) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p st.shared.f32 [%0], %1;\n\t}" ::"r"(addr), "f"(v),
               "r"((uint32_t)pred)
               : "memory");
}
// One tile's probabilities, staged densely ([rows][C] floats at `stage`), to global memory with 16-byte
// stores where alignment allows: `valid` floats, 128 cooperating threads (tid 0..127).
__device__ __forceinline__ void store_tile_probs(float* o, const float* stage, int valid, int tid) {
  if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) {
    const int nvec = valid >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(stage);
    float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
    for (int k = 0; k < 4; ++k) {            // 128 rows x 16 classes at most = 512 vectors
      const int i = tid + 128 * k;
      if (i < nvec) o4[i] = s4[i];
    }
    const int i = (nvec << 2) + tid;
    if (i < valid) o[i] = stage[i];
  } else {
    for (int i = tid; i < valid; i += 128) o[i] = stage[i];
  }
}

}  // namespace fusedk
