// Kernels 1+2+3 in one launch: uint8 queries in, ensemble-mean class probabilities out.
//
// Replaces, for trials that share a GPU, the whole of
//   InferenceWorker predict  (rafiki/worker/inference.py:53)
//     -> TfFeedForward.predict (examples/models/image_classification/TfFeedForward.py:89-97,
//        graph of :141-157: cast, Flatten, BatchNormalization, Dense+ReLU x L, Dense+softmax)
//     -> ensemble_predictions (rafiki/predictor/ensemble.py:10-14: mean over trials).
//
// Why one kernel.  A TfFeedForward trial has at most 128 hidden units (TfFeedForward.py:24),
// so each GEMM on its own is HBM-bound (arithmetic intensity ~ H flop/byte).  Keeping a
// 128-query tile on chip through all layers of all trials cuts HBM traffic to the
// algorithmic minimum, d bytes in + 4*C bytes out per query, at which point the work
// is bound by the tensor pipe (236 kflop per 824 bytes for D 784, H 128, L 2).
//
// Structure (one persistent CTA per SM, 352 threads, warp-specialised):
//   warps 0-3  epilogue   TMEM -> registers, bias + ReLU -> bf16 hidden tile in shared memory
//                         (the next layer's A operand); last layer: bias + softmax, and the
//                         running ensemble sum in float64 registers; final store of the mean
//   warps 4-7  A producer 16-byte loads of the uint8 tile, exact uint8 -> bf16 conversion,
//                         stores in the K-major 128B-swizzled layout tcgen05.mma reads
//   warp 8     W1 loader  cp.async.bulk (TMA engine) of the pre-swizzled first-layer weight
//                         blocks into a 4-stage ring shared with the A producer
//   warp 9     MMA issuer one thread issues every tcgen05.mma; accumulators live in TMEM
//                         (2 x 128 columns for layer 1, double buffered so the epilogue of one
//                         item overlaps the MMAs of the next, 128 for layer 2, 16 for layer 3)
//   warp 10    parameter loader  cp.async.bulk of W2, W3 and the biases of the item's trial
// A work item is (tile of 128 queries, trial); items of one tile run back to back so the tile
// stays hot in L1/L2 and the ensemble sum never leaves registers.
//
// BatchNormalization is folded into the first layer at pack time (mlp.cu), so the A operand
// is the raw pixel value, which bf16 holds exactly.
#include "pack.cuh"
#include "rfk_internal.cuh"
#include "sm100.cuh"

namespace {

using namespace sm100;

constexpr int TILE_M = 128;
constexpr int KB = 64;          // K elements per pipeline stage (one 128-byte swizzle atom of bf16)
constexpr int STAGES = 4;
constexpr int HMAX = 128;       // widest hidden layer the fused path takes
constexpr int CPAD = 16;        // class dimension padded to one UMMA N step
constexpr int A_STAGE = TILE_M * 128;
constexpr int B_STAGE = HMAX * 128;
constexpr int HID_ATOM = TILE_M * 128;
constexpr int HID_BYTES = 2 * HID_ATOM;
constexpr int W2_BYTES = 2 * HMAX * 128;
constexpr int W3_BYTES = 2 * CPAD * 128;
constexpr int BIAS_FLOATS = HMAX + HMAX + CPAD;   // b1 | b2 | b3
constexpr int BIAS_BYTES = BIAS_FLOATS * 4;

constexpr int OFF_A = 0;
constexpr int OFF_B = OFF_A + STAGES * A_STAGE;
constexpr int OFF_HID = OFF_B + STAGES * B_STAGE;
constexpr int OFF_W2 = OFF_HID + HID_BYTES;
constexpr int OFF_W3 = OFF_W2 + W2_BYTES;
constexpr int OFF_BIAS = OFF_W3 + W3_BYTES;
constexpr int OFF_BAR = OFF_BIAS + 2 * BIAS_BYTES;
constexpr int NUM_BARS = 2 * STAGES + 13;
constexpr int OFF_TMEM_SLOT = OFF_BAR + NUM_BARS * 8;
constexpr int SMEM_USED = OFF_TMEM_SLOT + 16;
constexpr int SMEM_BYTES = SMEM_USED + 1024;      // slack for the 1024-byte alignment of the base

constexpr int kEpiWarps = 4, kProdWarps = 4;
constexpr int kProdThreads = kProdWarps * 32;
constexpr int kThreads = (kEpiWarps + kProdWarps + 3) * 32;   // 352
constexpr int WARP_W1 = 8, WARP_MMA = 9, WARP_PAR = 10;

constexpr int TMEM_COLS = 512;
constexpr int COL_ACC1 = 0;     // two buffers of 128 columns
constexpr int COL_ACC2 = 256;
constexpr int COL_ACC3 = 384;

// when, during the layer-1 K loop of item i, the MMA thread slips in layers 2 and 3 of item i-1
constexpr int KB_SLOT_A = 3, KB_SLOT_B = 8;

struct TrialDev {
  const uint8_t* w1;    // [n_kb][h1_pad] rows of 128 B, swizzled
  const uint8_t* w2;    // [h1_pad/64 atoms][h2_pad] rows of 128 B (n_hidden == 2), else unused
  const uint8_t* w3;    // [hl_pad/64 atoms][16] rows of 128 B
  const float* bias;    // [BIAS_FLOATS]
  int h1_pad, h2_pad, n_hidden, pad_;
};

struct FusedParams {
  TrialDev trial[RFK_MAX_TRIALS];
  const uint8_t* x;
  float* out;
  long long n;
  int d, n_kb, w, total_trials, n_classes, n_tiles;
};

struct Bars {
  uint64_t full[STAGES], empty[STAGES];
  uint64_t acc1_full[2], acc1_empty[2];
  uint64_t hid_ready, acc2_full, acc3_full, w23_full, w23_free;
  uint64_t bias_full[2], bias_free[2];
};
static_assert(sizeof(Bars) == NUM_BARS * 8, "barrier block size");

// exact uint8 -> bf16: float(b) = as_float(0x4B000000 | b) - 2^23 has at most 8 significant
// bits, so its upper 16 bits ARE the bf16 value (no rounding).
__device__ __forceinline__ uint32_t u8x2_to_bf16x2(uint32_t word, int e0) {
  const float f0 = __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | e0)) - 8388608.0f;
  const float f1 = __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | (e0 + 1))) - 8388608.0f;
  return __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632u);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

__device__ __forceinline__ uint4 ldg_u4(const uint8_t* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__device__ __forceinline__ void sts_u4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// TMEM accumulator row -> bias + ReLU -> bf16 -> hidden tile (A operand of the next layer)
__device__ __forceinline__ void epilogue_hidden(uint32_t tmem_row_addr, int width, const float* bias,
                                                uint32_t hid_base, int row) {
  const uint32_t row_off = (uint32_t)row * 128u;
  const uint32_t sw = (uint32_t)(row & 7);
  for (int c0 = 0; c0 < width; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tmem_row_addr + (uint32_t)c0, v);
    tmem_ld_wait();
    uint32_t p[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float a = fmaxf(__uint_as_float(v[2 * j]) + bias[c0 + 2 * j], 0.f);
      const float b = fmaxf(__uint_as_float(v[2 * j + 1]) + bias[c0 + 2 * j + 1], 0.f);
      p[j] = pack_bf16x2(a, b);
    }
    const uint32_t atom = hid_base + (uint32_t)(c0 >> 6) * HID_ATOM + row_off;
    const uint32_t ch = (uint32_t)((c0 & 63) >> 3);
    sts_u4(atom + (((ch) ^ sw) << 4), make_uint4(p[0], p[1], p[2], p[3]));
    sts_u4(atom + (((ch + 1) ^ sw) << 4), make_uint4(p[4], p[5], p[6], p[7]));
  }
}

__global__ void __launch_bounds__(kThreads, 1)
fused_ensemble_kernel(const __grid_constant__ FusedParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  Bars* bars = reinterpret_cast<Bars*>(smem + OFF_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_TMEM_SLOT);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bars->full[s], kProdThreads + 1);
      mbar_init(&bars->empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->acc1_full[b], 1);
      mbar_init(&bars->acc1_empty[b], kEpiWarps * 32);
      mbar_init(&bars->bias_full[b], 1);
      mbar_init(&bars->bias_free[b], kEpiWarps * 32);
    }
    mbar_init(&bars->hid_ready, kEpiWarps * 32);
    mbar_init(&bars->acc2_full, 1);
    mbar_init(&bars->acc3_full, 1);
    mbar_init(&bars->w23_full, 1);
    mbar_init(&bars->w23_free, 1);
    fence_mbar_init();
  }
  if (warp == WARP_MMA) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  const int n_kb = P.n_kb;
  const int w = P.w;
  const uint32_t smem_base = smem_u32(smem);

  if (warp < kEpiWarps) {
    // =============================================================== epilogue warps
    const int row = threadIdx.x;                               // TMEM lane == row of the tile
    const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
    const uint32_t hid_base = smem_base + OFF_HID;
    double acc[CPAD];
#pragma unroll
    for (int j = 0; j < CPAD; ++j) acc[j] = 0.0;
    int item = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      for (int t = 0; t < w; ++t, ++item) {
        const TrialDev& T = P.trial[t];
        const int buf = item & 1;
        const uint32_t par2 = (uint32_t)((item >> 1) & 1);
        const uint32_t par1 = (uint32_t)(item & 1);
        const float* bias = reinterpret_cast<const float*>(smem + OFF_BIAS + buf * BIAS_BYTES);
        mbar_wait(&bars->bias_full[buf], par2);
        // ---- layer 1: acc1[buf] -> hidden
        mbar_wait(&bars->acc1_full[buf], par2);
        tc_fence_after_sync();
        epilogue_hidden(lane_addr + COL_ACC1 + buf * HMAX, T.h1_pad, bias, hid_base, row);
        tc_fence_before_sync();
        mbar_arrive(&bars->acc1_empty[buf]);
        fence_proxy_async_smem();
        mbar_arrive(&bars->hid_ready);
        // ---- layer 2 (two-hidden-layer trials): acc2 -> hidden
        if (T.n_hidden == 2) {
          mbar_wait(&bars->acc2_full, par1);
          tc_fence_after_sync();
          epilogue_hidden(lane_addr + COL_ACC2, T.h2_pad, bias + HMAX, hid_base, row);
          tc_fence_before_sync();
          fence_proxy_async_smem();
          mbar_arrive(&bars->hid_ready);
        }
        // ---- softmax layer: acc3 -> probabilities -> running ensemble sum
        mbar_wait(&bars->acc3_full, par1);
        tc_fence_after_sync();
        uint32_t v[16];
        tmem_ld16(lane_addr + COL_ACC3, v);
        tmem_ld_wait();
        tc_fence_before_sync();
        float z[CPAD];
        float mx = -INFINITY;
#pragma unroll
        for (int j = 0; j < CPAD; ++j) {
          z[j] = __uint_as_float(v[j]) + bias[2 * HMAX + j];
          if (j < P.n_classes) mx = fmaxf(mx, z[j]);
        }
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < CPAD; ++j) {
          z[j] = j < P.n_classes ? expf(z[j] - mx) : 0.f;
          sum += z[j];
        }
        mbar_arrive(&bars->bias_free[buf]);
#pragma unroll
        for (int j = 0; j < CPAD; ++j) acc[j] += (double)(z[j] / sum);
      }
      // ---- the tile's ensemble mean: float64 sum in trial order, one division, one rounding
      const long long grow = (long long)tile * TILE_M + row;
      if (grow < P.n) {
        float* o = P.out + grow * P.n_classes;
#pragma unroll
        for (int j = 0; j < CPAD; ++j)
          if (j < P.n_classes) o[j] = (float)(acc[j] / (double)P.total_trials);
      }
#pragma unroll
      for (int j = 0; j < CPAD; ++j) acc[j] = 0.0;
    }
  } else if (warp < kEpiWarps + kProdWarps) {
    // =============================================================== A producer warps
    const int tp = threadIdx.x - kEpiWarps * 32;                 // 0..127
    const int q = tp & 3;                                        // 16-byte quarter of the 64-byte row chunk
    const int r_in = tp >> 2;                                    // 0..31
    const uint32_t a_base = smem_base + OFF_A;
    int stage = 0;
    uint32_t phase = 0;
    const int d = P.d;
    // register double buffering: the loads of step s+1 are in flight while step s is converted
    uint4 cur[4], nxt[4];
    auto load_step = [&](int tile, int kb, uint4 (&dst)[4]) {
      const long long row0 = (long long)tile * TILE_M;
      const int k = kb * KB + q * 16;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const long long grow = row0 + it * 32 + r_in;
        dst[it] = (grow < P.n && k < d) ? ldg_u4(P.x + grow * d + k) : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    // flattened step cursor for the prefetch: (tile, t, kb)
    int l_tile = blockIdx.x, l_t = 0, l_kb = 0;
    auto advance_load = [&]() {
      if (++l_kb == n_kb) {
        l_kb = 0;
        if (++l_t == w) {
          l_t = 0;
          l_tile += gridDim.x;
        }
      }
    };
    if (l_tile < P.n_tiles) load_step(l_tile, l_kb, cur);
    advance_load();
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      for (int t = 0; t < w; ++t) {
        for (int kb = 0; kb < n_kb; ++kb) {
          if (l_tile < P.n_tiles) load_step(l_tile, l_kb, nxt);
          advance_load();
          mbar_wait(&bars->empty[stage], phase ^ 1u);
          const uint32_t st_base = a_base + (uint32_t)stage * A_STAGE;
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            const int row = it * 32 + r_in;
            const uint4 raw = cur[it];
            const uint4 lo = make_uint4(u8x2_to_bf16x2(raw.x, 0), u8x2_to_bf16x2(raw.x, 2),
                                        u8x2_to_bf16x2(raw.y, 0), u8x2_to_bf16x2(raw.y, 2));
            const uint4 hi = make_uint4(u8x2_to_bf16x2(raw.z, 0), u8x2_to_bf16x2(raw.z, 2),
                                        u8x2_to_bf16x2(raw.w, 0), u8x2_to_bf16x2(raw.w, 2));
            const uint32_t rbase = st_base + (uint32_t)row * 128u;
            const uint32_t sw = (uint32_t)(row & 7);
            sts_u4(rbase + ((((uint32_t)(2 * q)) ^ sw) << 4), lo);
            sts_u4(rbase + ((((uint32_t)(2 * q + 1)) ^ sw) << 4), hi);
          }
          fence_proxy_async_smem();
          mbar_arrive(&bars->full[stage]);
#pragma unroll
          for (int it = 0; it < 4; ++it) cur[it] = nxt[it];
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == WARP_W1) {
    // =============================================================== first-layer weight loader
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        for (int t = 0; t < w; ++t) {
          const TrialDev& T = P.trial[t];
          const uint32_t bytes = (uint32_t)T.h1_pad * 128u;
          for (int kb = 0; kb < n_kb; ++kb) {
            mbar_wait(&bars->empty[stage], phase ^ 1u);
            mbar_arrive_expect_tx(&bars->full[stage], bytes);
            bulk_g2s(smem + OFF_B + stage * B_STAGE, T.w1 + (size_t)kb * bytes, bytes, &bars->full[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == WARP_PAR) {
    // =============================================================== per-trial parameter loader
    if (lane == 0) {
      int item = 0;
      for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        for (int t = 0; t < w; ++t, ++item) {
          const TrialDev& T = P.trial[t];
          const int buf = item & 1;
          mbar_wait(&bars->bias_free[buf], (uint32_t)(((item >> 1) & 1) ^ 1));
          mbar_arrive_expect_tx(&bars->bias_full[buf], BIAS_BYTES);
          bulk_g2s(smem + OFF_BIAS + buf * BIAS_BYTES, T.bias, BIAS_BYTES, &bars->bias_full[buf]);
          mbar_wait(&bars->w23_free, (uint32_t)((item & 1) ^ 1));
          const int hl_pad = T.n_hidden == 2 ? T.h2_pad : T.h1_pad;
          const uint32_t w3_bytes = (uint32_t)((hl_pad + 63) / 64) * CPAD * 128u;
          const uint32_t w2_bytes = T.n_hidden == 2 ? (uint32_t)((T.h1_pad + 63) / 64) * (uint32_t)T.h2_pad * 128u : 0u;
          mbar_arrive_expect_tx(&bars->w23_full, w2_bytes + w3_bytes);
          if (w2_bytes) bulk_g2s(smem + OFF_W2, T.w2, w2_bytes, &bars->w23_full);
          bulk_g2s(smem + OFF_W3, T.w3, w3_bytes, &bars->w23_full);
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // =============================================================== MMA issuer (one thread)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0, hid_phase = 0;
      const uint32_t a_base = smem_base + OFF_A, b_base = smem_base + OFF_B;
      const uint32_t hid_base = smem_base + OFF_HID, w2_base = smem_base + OFF_W2, w3_base = smem_base + OFF_W3;
      // layers 2/3 of a finished layer-1 item still to be issued: pend 1 = layer 2 next, 2 = layer 3 next
      int pend = 0, pend_item = 0, pend_t = 0;
      auto issue_layer2 = [&]() {
        const TrialDev& T = P.trial[pend_t];
        mbar_wait(&bars->w23_full, (uint32_t)(pend_item & 1));
        mbar_wait(&bars->hid_ready, hid_phase);
        hid_phase ^= 1u;
        tc_fence_after_sync();
        const uint32_t idesc = idesc_bf16_f32(TILE_M, T.h2_pad);
        const int ksteps = T.h1_pad / 16;
        for (int k = 0; k < ksteps; ++k) {
          const uint32_t a = hid_base + (uint32_t)(k >> 2) * HID_ATOM + (uint32_t)(k & 3) * 32u;
          const uint32_t b = w2_base + (uint32_t)(k >> 2) * ((uint32_t)T.h2_pad * 128u) + (uint32_t)(k & 3) * 32u;
          mma_bf16_ss(tmem_base + COL_ACC2, smem_desc_sw128(a), smem_desc_sw128(b), idesc, k > 0);
        }
        mma_commit(&bars->acc2_full);
        pend = 2;
      };
      auto issue_layer3 = [&]() {
        const TrialDev& T = P.trial[pend_t];
        if (T.n_hidden != 2) mbar_wait(&bars->w23_full, (uint32_t)(pend_item & 1));
        mbar_wait(&bars->hid_ready, hid_phase);
        hid_phase ^= 1u;
        tc_fence_after_sync();
        const uint32_t idesc = idesc_bf16_f32(TILE_M, CPAD);
        const int hl_pad = T.n_hidden == 2 ? T.h2_pad : T.h1_pad;
        const int ksteps = hl_pad / 16;
        for (int k = 0; k < ksteps; ++k) {
          const uint32_t a = hid_base + (uint32_t)(k >> 2) * HID_ATOM + (uint32_t)(k & 3) * 32u;
          const uint32_t b = w3_base + (uint32_t)(k >> 2) * (CPAD * 128u) + (uint32_t)(k & 3) * 32u;
          mma_bf16_ss(tmem_base + COL_ACC3, smem_desc_sw128(a), smem_desc_sw128(b), idesc, k > 0);
        }
        mma_commit(&bars->acc3_full);
        mma_commit(&bars->w23_free);
        pend = 0;
      };
      auto issue_pending = [&]() {
        if (pend == 1) issue_layer2();
        else if (pend == 2) issue_layer3();
      };
      int item = 0;
      for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        for (int t = 0; t < w; ++t, ++item) {
          const TrialDev& T = P.trial[t];
          const int buf = item & 1;
          const uint32_t idesc1 = idesc_bf16_f32(TILE_M, T.h1_pad);
          const uint32_t d_tmem = tmem_base + COL_ACC1 + buf * HMAX;
          mbar_wait(&bars->acc1_empty[buf], (uint32_t)(((item >> 1) & 1) ^ 1));
          tc_fence_after_sync();
          for (int kb = 0; kb < n_kb; ++kb) {
            mbar_wait(&bars->full[stage], phase);
            tc_fence_after_sync();
            const int rem = (P.d - kb * KB) / 16;
            const int ksteps = rem < 4 ? rem : 4;
            const uint32_t a = a_base + (uint32_t)stage * A_STAGE;
            const uint32_t b = b_base + (uint32_t)stage * B_STAGE;
            for (int k = 0; k < ksteps; ++k)
              mma_bf16_ss(d_tmem, smem_desc_sw128(a + k * 32), smem_desc_sw128(b + k * 32), idesc1, (kb | k) != 0);
            mma_commit(&bars->empty[stage]);
            if (kb == n_kb - 1) mma_commit(&bars->acc1_full[buf]);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            if (kb == KB_SLOT_A || kb == KB_SLOT_B) issue_pending();
          }
          while (pend != 0) issue_pending();        // short K loops: finish the previous item here
          pend = T.n_hidden == 2 ? 1 : 2;
          pend_item = item;
          pend_t = t;
        }
      }
      while (pend != 0) issue_pending();
    }
  }
  __syncwarp();
  tc_fence_before_sync();
  __syncthreads();
  if (warp == WARP_MMA) tmem_dealloc<TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------- host side
struct FusedState {
  uint8_t* w1 = nullptr;
  uint8_t* w2 = nullptr;
  uint8_t* w3 = nullptr;
  float* bias = nullptr;
  TrialDev dev;
};

template <typename T>
int upload_vec(T** dst, const std::vector<T>& v) {
  RFK_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(dst), v.size() * sizeof(T)));
  RFK_CUDA_TRY(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return RFK_OK;
}

}  // namespace

namespace rfk {

bool fused_applicable(const rfk_mlp* h) {
  if (h->dtype != RFK_BF16) return false;
  if (h->n_hidden < 1 || h->n_hidden > 2) return false;
  if (h->n_classes > CPAD || h->n_classes > RFK_MAX_CLASSES) return false;
  if (h->d % 16 != 0) return false;
  for (int i = 0; i < h->n_hidden; ++i)
    if (h->layers[i].out > HMAX) return false;
  return true;
}

int fused_prepare(rfk_mlp* h) {
  FusedState* s = new FusedState();
  h->fused = s;
  const Layer& l1 = h->layers[0];
  const Layer& l3 = h->layers[h->n_hidden];
  std::vector<uint8_t> img;
  pack_kmajor_sw128(l1.w_bf16_host.data(), l1.out_pad, l1.in_pad, &img);
  int rc = upload_vec(&s->w1, img);
  if (rc != RFK_OK) return rc;
  std::vector<float> bias(BIAS_FLOATS, 0.f);
  for (int j = 0; j < l1.out_pad; ++j) bias[j] = l1.b_fold_host[j];
  if (h->n_hidden == 2) {
    const Layer& l2 = h->layers[1];
    pack_kmajor_sw128(l2.w_bf16_host.data(), l2.out_pad, l2.in_pad, &img);
    rc = upload_vec(&s->w2, img);
    if (rc != RFK_OK) return rc;
    for (int j = 0; j < l2.out_pad; ++j) bias[HMAX + j] = l2.b_fold_host[j];
  }
  pack_kmajor_sw128(l3.w_bf16_host.data(), l3.out_pad, l3.in_pad, &img);
  rc = upload_vec(&s->w3, img);
  if (rc != RFK_OK) return rc;
  for (int j = 0; j < l3.out_pad && j < CPAD; ++j) bias[2 * HMAX + j] = l3.b_fold_host[j];
  rc = upload_vec(&s->bias, bias);
  if (rc != RFK_OK) return rc;
  s->dev.w1 = s->w1;
  s->dev.w2 = s->w2;
  s->dev.w3 = s->w3;
  s->dev.bias = s->bias;
  s->dev.h1_pad = l1.out_pad;
  s->dev.h2_pad = h->n_hidden == 2 ? h->layers[1].out_pad : 0;
  s->dev.n_hidden = h->n_hidden;
  s->dev.pad_ = 0;
  static bool attr_set[64] = {false};
  if (!attr_set[h->device & 63]) {
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_ensemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_set[h->device & 63] = true;
  }
  return RFK_OK;
}

void fused_release(rfk_mlp* h) {
  FusedState* s = static_cast<FusedState*>(h->fused);
  if (s == nullptr) return;
  cudaFree(s->w1);
  cudaFree(s->w2);
  cudaFree(s->w3);
  cudaFree(s->bias);
  delete s;
  h->fused = nullptr;
}

int fused_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n, float* out,
                  cudaStream_t st) {
  RFK_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "fused forward: x must be 16-byte aligned");
  FusedParams P;
  memset(&P, 0, sizeof(P));
  for (int t = 0; t < w; ++t) {
    const FusedState* s = static_cast<const FusedState*>(trials[t]->fused);
    RFK_REQUIRE(s != nullptr, "fused forward: trial %d was not packed for the fused path", t);
    P.trial[t] = s->dev;
  }
  P.x = x;
  P.out = out;
  P.n = n;
  P.d = trials[0]->d;
  P.n_kb = (P.d + KB - 1) / KB;
  P.w = w;
  P.total_trials = total_trials;
  P.n_classes = trials[0]->n_classes;
  P.n_tiles = (int)((n + TILE_M - 1) / TILE_M);
  const int sms = num_sms();
  const int grid = P.n_tiles < sms ? P.n_tiles : sms;
  fused_ensemble_kernel<<<grid, kThreads, SMEM_BYTES, st>>>(P);
  RFK_LAUNCH_CHECK("fused_ensemble_kernel");
  return RFK_OK;
}

}  // namespace rfk
