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
// Structure (one persistent CTA per SM, 640 threads in five warpgroups, warp-specialised; the
// epilogue warpgroups take registers the light roles release, setmaxnreg):
//   warps 0-7   epilogue   TMEM -> registers, bias + ReLU -> 16-bit hidden activations, written back
//                          to tensor memory (the next layer's A operand: no shared-memory round
//                          trip).  Warps 0-3 own hidden columns 0-63, warps 4-7 columns 64-127 (a
//                          warp may only touch the TMEM lanes 32*(warp%4)..+31, so two warps share
//                          each lane quarter).  Warps 0-3 also do the last layer: bias + softmax, the
//                          running ensemble sum in float64 registers, and the final store of the mean.
//   warps 8-15  converter  raw uint8 tile (shared memory, written by TMA with the 128-byte swizzle)
//                          -> fp16 -> the layer-1 A operand, stored straight into TENSOR memory
//                          (tcgen05.st): the converted tile never touches shared memory.  With fp16
//                          operands the conversion is a zero-extension: the pattern 0x00bb is the
//                          subnormal b * 2^-24, and the 2^24 is folded into the layer-1 bias FFMA.
//   warp 16     loader     per 128-wide K step one cp.async.bulk of the pre-swizzled first-layer
//                          weight block into a 4-deep ring and one TMA tensor load of the 128 x 128-byte
//                          uint8 tile (out-of-range rows / columns arrive as zeros) into a 3-deep
//                          raw ring; W2, W3 and the biases of the next items as a third cursor
//   warp 17     MMA issuer, layer 1: A from tensor memory, B from shared memory; accumulators in
//                          TMEM (2 x 128 columns, double buffered so the epilogue of one item
//                          overlaps the MMAs of the next).  ONE thread issues every MMA of an
//                          accumulator: tcgen05.mma instructions of one thread execute in issue
//                          order, those of two threads do not (measured: two warps taking alternate
//                          K steps behind a named-barrier token gave results that differed in the last
//                          bits from launch to launch, because fp32 accumulation is not associative).
//   warp 18     spare (keeps warpgroup 4 complete for setmaxnreg)
//   warp 19     MMA issuer for layers 2 and 3 (hidden tile x W2 / W3); layer 2 accumulates into the
//                          layer-1 buffer its epilogue has just drained, layer 3 into 16 columns
// Every tcgen05.mma / TMA / bulk-copy issue is warp-convergent with an elect.sync predicate on the
// instruction itself, and the warp index is taken through a shuffle so that the compiler knows it is
// warp-uniform: addresses, descriptors and loop counters of the issuing roles then live in uniform
// registers and each MMA / bulk copy is one instruction.  (With `threadIdx.x >> 5` every operand
// went through R2UR moves, ~60 cycles per MMA; from inside an `if (lane == 0)` region each issue
// costs a few hundred cycles.)
// A work item is (tile of 128 queries, trial); items of one tile run back to back so the tile
// stays hot in L2 and the ensemble sum never leaves registers.
//
// BatchNormalization is folded into the first layer at pack time (mlp.cu), so the A operand
// is the raw pixel value, which fp16 and bf16 both hold exactly.  Operands are fp16 (RFK_F16,
// the default: 11-bit significands keep the probabilities within 2e-3 of the fp32 oracle) or
// bf16 (RFK_BF16); accumulation is fp32 in TMEM either way.
#include <stdlib.h>

#include <atomic>

#include "pack.cuh"
#include "rfk_internal.cuh"
#include "sm100.cuh"
#include "tmap.cuh"
#include "fused_common.cuh"

namespace {

using namespace sm100;
using namespace fusedk;

constexpr int TILE_M = 128;
constexpr int ATOM_K = 64;      // K elements of one 128-byte swizzle atom of 16-bit operands
constexpr int KB = 128;         // K elements per pipeline step: two atoms, so that every barrier round
                                // (loader -> converter -> MMA issuer) carries eight MMAs instead of four
constexpr int A_STAGES = 2;     // A operand ring: converted query tile, in TENSOR memory (64 columns per stage)
constexpr int B_STAGES = 4;     // B operand ring: first-layer weight blocks in shared memory; deep, because a
                                // 32 KB block takes 1000-1500 cycles to arrive from L2 under load.  (Eight
                                // stages of one 16 KB atom each, handed back after four MMAs, measured 17 %
                                // slower: the copies, not their latency, are what the ring waits for.)
constexpr int RAW_STAGES = 3;   // raw uint8 ring (TMA landing zone; covers HBM latency)
constexpr int HMAX = 128;       // widest hidden layer the fused path takes
constexpr int CPAD = 16;        // class dimension padded to one UMMA N step
constexpr int B_STAGE = (KB / ATOM_K) * HMAX * 128;
constexpr int RAW_STAGE = TILE_M * KB;            // 8192 bytes of uint8
constexpr int W2_BYTES = 2 * HMAX * 128;
constexpr int W3_BYTES = 2 * CPAD * 128;
constexpr int BIAS_FLOATS = HMAX + HMAX + CPAD;   // b1 | b2 | b3
constexpr int BIAS_BYTES = BIAS_FLOATS * 4;

constexpr int OFF_B = 0;
constexpr int OFF_RAW = OFF_B + B_STAGES * B_STAGE;
constexpr int OFF_W2 = OFF_RAW + RAW_STAGES * RAW_STAGE;
constexpr int OFF_W3 = OFF_W2 + W2_BYTES;
constexpr int OFF_BIAS = OFF_W3 + 2 * W3_BYTES;     // W3 is double buffered by item parity
constexpr int OUT_BYTES = TILE_M * CPAD * 4;       // staging of one tile's probabilities for 16-byte stores
constexpr int OFF_OUT = (OFF_BIAS + 2 * BIAS_BYTES + 15) / 16 * 16;
constexpr int OFF_BAR = OFF_OUT + OUT_BYTES;
constexpr int NUM_BARS = A_STAGES + 2 * B_STAGES + 2 * RAW_STAGES + 17;
constexpr int OFF_TMEM_SLOT = OFF_BAR + NUM_BARS * 8;
constexpr int SMEM_USED = OFF_TMEM_SLOT + 16;
constexpr int SMEM_BYTES = SMEM_USED;             // the kernel has no static shared memory, so the dynamic
                                                  // window starts 1024-byte aligned (checked at kernel entry)
static_assert(OFF_W2 % 1024 == 0 && OFF_W3 % 1024 == 0 && OFF_B % 1024 == 0 && OFF_RAW % 1024 == 0, "operand alignment");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

// Five warpgroups (setmaxnreg works per warpgroup): 0-1 epilogue, 2-3 converters, 4 = loader, MMA
// issuer for layer 1, a spare warp, MMA issuer for layers 2/3.  The kernel starts with 96 registers per thread;
// the light roles give registers back and the epilogue warps take them.
constexpr int kEpiWarps = 8, kConvWarps = 8;
constexpr int kEpiThreads = kEpiWarps * 32;                   // 256
constexpr int kConvThreads = kConvWarps * 32;                 // 256: half a row (64 bytes) per thread per step
constexpr int kThreads = (kEpiWarps + kConvWarps + 4) * 32;   // 640
constexpr int WARP_LOAD = kEpiWarps + kConvWarps, WARP_MMA = WARP_LOAD + 1, WARP_TAIL = WARP_LOAD + 3;
constexpr int kRegsLaunch = 96;   // registers per thread the kernel is compiled for and launched with
constexpr int kRegsEpi = 136, kRegsConv = 72, kRegsLight = 56;
// setmaxnreg.inc only hands out what the CTA's own warps have given back: the pool is the launch
// allocation, not the register file (an over-subscribed budget leaves one epilogue warp blocked for ever)
static_assert(kEpiWarps * kRegsEpi + kConvWarps * kRegsConv + 4 * kRegsLight <= (kEpiWarps + kConvWarps + 4) * kRegsLaunch,
              "setmaxnreg budget exceeds the launch allocation");
// named barriers (id 0 is __syncthreads): the four softmax warps; converters -> layer-1 issuer, one per A stage;
// epilogue warps -> layers-2/3 issuer
constexpr int BAR_SOFTMAX = 1, BAR_A0 = 2 /* .. BAR_A0 + A_STAGES - 1 */, BAR_HID = BAR_A0 + 4;

constexpr int TMEM_COLS = 512;
constexpr int COL_ACC1 = 0;     // two buffers of 128 columns; layer 2 accumulates into the buffer its own
                                // layer-1 epilogue has just drained
constexpr int COL_A = 256;      // A operand ring of layer 1: A_STAGES x 64 columns of 16-bit pairs
constexpr int COL_HID = 384;    // hidden activations, 16-bit pairs: 64 columns (the next layer's A operand)
constexpr int COL_ACC3 = 448;
static_assert(COL_A + A_STAGES * (KB / 2) <= COL_HID && COL_ACC3 + CPAD <= TMEM_COLS, "tensor memory map");

struct TrialDev {
  const uint8_t* w1;    // [n_kb][h1_pad] rows of 128 B, swizzled
  const uint8_t* w2;    // [h1_pad/64 atoms][h2_pad] rows of 128 B (n_hidden == 2), else unused
  const uint8_t* w3;    // [hl_pad/64 atoms][16] rows of 128 B
  const float* bias;    // [BIAS_FLOATS]
  int h1_pad, h2_pad, n_hidden, pad_;
};

struct FusedParams {
  TrialDev trial[RFK_MAX_TRIALS];
  float* out;
  long long n;
  int d, n_kb, w, total_trials, n_classes, n_tiles;
  long long* prof;   // development only (RFK_FUSED_PROF): PF_COUNT counters, summed over CTAs
  long long* trace;  // development only (RFK_FUSED_PROF=2): block 0 event timestamps, [5 roles][kTraceLen]
  int split;         // 1: one (tile, trial) item per CTA, per-trial probabilities to out[t][n][C] (small batches)
  int debug;         // development only (RFK_FUSED_DEBUG): 1 skip W1 copies, 2 skip conversion, 4 skip layer-1 MMAs
};

struct Bars {
  uint64_t a_full[A_STAGES];                      // converters -> MMA issuer
  uint64_t b_full[B_STAGES];                      // weight block landed
  uint64_t step_done[B_STAGES];                   // MMAs of a step complete: frees its B stage and its A stage
  uint64_t raw_full[RAW_STAGES], raw_empty[RAW_STAGES];
  uint64_t acc1_full[2], acc1_empty[2];
  uint64_t hid_ready, acc2_full, acc3_full;
  uint64_t w2_full, w2_free, w3_full[2], w3_free[2];
  uint64_t bias_full[2], bias_free[2];
};
static_assert(sizeof(Bars) == NUM_BARS * 8, "barrier block size");


// kProf (development, RFK_FUSED_PROF=1): the leader thread of each warp role accumulates the
// cycles it spends blocked on each kind of barrier and adds them to P.prof at exit.
enum ProfSlot { PF_EPI_BIAS, PF_EPI_ACC1, PF_EPI_ACC2, PF_EPI_ACC3, PF_EPI_TOTAL, PF_CONV_RAW, PF_CONV_EMPTY,
                PF_CONV_TOTAL, PF_LOAD_EMPTY, PF_LOAD_RAW_EMPTY, PF_PAR_BIAS_FREE, PF_PAR_W23_FREE, PF_MMA_ACC1_EMPTY,
                PF_MMA_FULL, PF_MMA_W23_FULL, PF_MMA_HID_L2, PF_MMA_HID_L3, PF_MMA_TOTAL, PF_COUNT };

constexpr int kTraceLen = 2048;
// append a timestamp to a role's trace (block 0, leader lane only)
#define RFK_TRACE(role, tag)                                                                      \
  do {                                                                                           \
    if (kProf && P.trace != nullptr && blockIdx.x == 0 && lane == 0 && trace_n < kTraceLen) {      \
      P.trace[(role) * kTraceLen + trace_n] = (clock64() << 4) | (long long)(tag);                \
      ++trace_n;                                                                                 \
    }                                                                                            \
  } while (0)

template <bool kProf>
__device__ __forceinline__ void wait_prof(uint64_t* bar, uint32_t parity, long long (&prof)[PF_COUNT], int slot) {
  if (kProf) {
    const long long t0 = clock64();
    mbar_wait(bar, parity);
    prof[slot] += clock64() - t0;
  } else {
    mbar_wait(bar, parity);
  }
}

template <bool kF16, bool kProf>
__global__ void __launch_bounds__(kThreads, 1)
fused_ensemble_kernel(const __grid_constant__ FusedParams P, const __grid_constant__ CUtensorMap x_map) {
  long long prof[PF_COUNT];
  if (kProf) {
#pragma unroll
    for (int i = 0; i < PF_COUNT; ++i) prof[i] = 0;
  }
  const long long t_start = kProf ? clock64() : 0;
  int trace_n = 0;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  Bars* bars = reinterpret_cast<Bars*>(smem + OFF_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_TMEM_SLOT);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  if ((smem_u32(smem) & 1023u) != 0u) {        // the swizzled operand tiles need 1024-byte alignment
    if (threadIdx.x == 0) printf("rafiki_b200: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }

  if (threadIdx.x == 0) {
    for (int s = 0; s < B_STAGES; ++s) {
      mbar_init(&bars->b_full[s], 1);
      mbar_init(&bars->step_done[s], 1);
    }
    for (int s = 0; s < RAW_STAGES; ++s) {
      mbar_init(&bars->raw_full[s], 1);
      mbar_init(&bars->raw_empty[s], kConvWarps);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->acc1_full[b], 1);
      mbar_init(&bars->acc1_empty[b], kEpiWarps);
      mbar_init(&bars->bias_full[b], 1);
      mbar_init(&bars->bias_free[b], kEpiWarps);
    }
    mbar_init(&bars->acc2_full, 1);
    mbar_init(&bars->acc3_full, 1);
    mbar_init(&bars->w2_full, 1);
    mbar_init(&bars->w2_free, 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->w3_full[b], 1);
      mbar_init(&bars->w3_free[b], 1);
    }
    fence_mbar_init();
  }
  if (warp == WARP_LOAD && lane == 0) tma_prefetch_desc(&x_map);
  if (warp == WARP_MMA) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  // The kernel owns all 512 TMEM columns of its SM (one CTA per SM), so the allocation starts at
  // column 0, lane 0.  Using the literal 0 keeps every TMEM address warp-uniform for the compiler.
  if (*tmem_slot != 0u) {
    if (threadIdx.x == 0) printf("rafiki_b200: unexpected TMEM base %u\n", *tmem_slot);
    __trap();
  }
  constexpr uint32_t tmem_base = 0u;
  // launched as a programmatic dependent of the previous kernel of the stream (see fused_pair.cu): the set-up
  // above overlapped that kernel's tail; nothing touches global memory before the previous grid is complete
  griddep_launch_dependents();
  griddep_wait();

  const int n_kb = P.n_kb;
  // Work assignment.  Normal: CTA b takes tiles b, b + grid, ... and runs all trials of a tile back
  // to back.  Split (small batches): CTA b is exactly one item, tile b / W, trial b % W, so that the
  // trials of a tile run on different SMs at the same time; kernel 3 then averages the W outputs.
  const int tile0 = P.split ? (int)blockIdx.x / P.w : (int)blockIdx.x;
  const int tile_step = P.split ? P.n_tiles : (int)gridDim.x;
  const int t0 = P.split ? (int)blockIdx.x % P.w : 0;
  const int t1 = P.split ? t0 + 1 : P.w;
  const int w = t1 - t0;                       // trials this CTA runs per tile
  const uint32_t smem_base = smem_u32(smem);

  if (warp < kEpiWarps) {
    // =============================================================== epilogue warps
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsEpi));
    const int quarter = warp & 3;                              // TMEM lane quarter this warp may touch
    const int half = warp >> 2;                                // 0: hidden columns 0-63 + softmax; 1: columns 64-127
    const int row = quarter * 32 + lane;                       // TMEM lane == row of the tile
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t hid_row_addr = lane_addr + COL_HID;
    double acc[CPAD];
#pragma unroll
    for (int j = 0; j < CPAD; ++j) acc[j] = 0.0;
    // The softmax of item i-1 is deferred until the layer-1 epilogue of item i has been published:
    // it then runs while the tensor pipe works on layer 2 of item i instead of sitting on the
    // item-to-item critical path.  Its accumulator row is pulled into registers first (v3), because
    // a one-hidden-layer item i may overwrite acc3 as soon as its hidden tile is published.
    uint32_t v3[16];
    bool have_prev = false;
    int prev_tile = 0, prev_t = 0, prev_buf = 0;
    auto finish_prev = [&]() {
      // half 0: logits (v3) + bias -> softmax -> running ensemble sum; store the mean after the last trial
      const float* pbias = reinterpret_cast<const float*>(smem + OFF_BIAS + prev_buf * BIAS_BYTES);
      if (half == 0) {
        float z[CPAD];
        const float inv = softmax16(v3, pbias + 2 * HMAX, P.n_classes, z);
        warp_arrive(&bars->bias_free[prev_buf]);
#pragma unroll
        for (int j = 0; j < CPAD; ++j) acc[j] += (double)(z[j] * inv);
        if (warp == 0) RFK_TRACE(3, 6);
        if (prev_t == t1 - 1) {
          // the tile's ensemble mean: float64 sum in trial order, one division, one rounding.  The
          // 128 rows of a tile are one contiguous block of the output, so they are staged in shared
          // memory and leave as 16-byte stores (the output may be peer memory across NVLink, where
          // 4-byte stores would waste most of every packet).
          const int C = P.n_classes;
          float* stage = reinterpret_cast<float*>(smem + OFF_OUT);
          const uint32_t dst = smem_base + OFF_OUT + (uint32_t)(row * C) * 4u;
          const double total = (double)(P.split ? 1 : P.total_trials);
#pragma unroll
          for (int j = 0; j < CPAD; ++j) sts_f32_if(dst + 4u * j, (float)(acc[j] / total), j < C);
#pragma unroll
          for (int j = 0; j < CPAD; ++j) acc[j] = 0.0;
          named_bar_sync(BAR_SOFTMAX, 128);                       // the four softmax warps
          {
            const long long first = (long long)prev_tile * TILE_M;
            const long long rows = P.n - first < TILE_M ? P.n - first : TILE_M;
            store_tile_probs(P.out + (P.split ? (long long)t0 * P.n * C : 0ll) + first * C, stage, (int)rows * C, row);
          }
          named_bar_sync(BAR_SOFTMAX, 128);                       // staging buffer free again
        }
      } else {
        warp_arrive(&bars->bias_free[prev_buf]);
      }
    };
    auto fetch_prev_logits = [&](uint32_t parity) {
      // both halves wait: the hidden tile may not be overwritten before the layer-3 MMAs have read it
      wait_prof<kProf>(&bars->acc3_full, parity, prof, PF_EPI_ACC3);
      if (warp == 0) RFK_TRACE(3, 5);
      tc_fence_after_sync();
      if (half == 0) {
        tmem_ld16(lane_addr + COL_ACC3, v3);
        tmem_ld_wait();
      }
    };
    int item = 0;
    uint32_t acc2_phase = 0;          // acc2_full completes one phase per two-hidden-layer item only
    if (P.debug & 64) {               // development: layer-1 pipeline alone (no epilogue work, no layers 2/3)
      for (int tile = tile0; tile < P.n_tiles; tile += tile_step)
        for (int t = t0; t < t1; ++t, ++item) {
          mbar_wait(&bars->acc1_full[item & 1], (uint32_t)((item >> 1) & 1));
          warp_arrive(&bars->acc1_empty[item & 1]);
        }
    }
    for (int tile = tile0; tile < ((P.debug & 64) ? 0 : P.n_tiles); tile += tile_step) {
      for (int t = t0; t < t1; ++t, ++item) {
        const TrialDev& T = P.trial[t];
        const int buf = item & 1;
        const uint32_t par2 = (uint32_t)((item >> 1) & 1);
        const float* bias = reinterpret_cast<const float*>(smem + OFF_BIAS + buf * BIAS_BYTES);
        if (w > 1 || item < 2) wait_prof<kProf>(&bars->bias_full[buf], par2, prof, PF_EPI_BIAS);
        // ---- layer 1: acc1[buf] -> hidden (after layer 3 of the previous item has read the hidden tile)
        wait_prof<kProf>(&bars->acc1_full[buf], par2, prof, PF_EPI_ACC1);
        if (warp == 0) RFK_TRACE(3, 1);
        if (have_prev) fetch_prev_logits((uint32_t)((item - 1) & 1));
        tc_fence_after_sync();
        {
          const int split = T.h1_pad < 64 ? T.h1_pad : 64;
          epilogue_hidden<kF16>(lane_addr + COL_ACC1 + buf * HMAX, half ? split : 0, half ? T.h1_pad : split, bias,
                                kF16 ? kSubnormalScale : 1.0f, hid_row_addr);
        }
        tc_fence_before_sync();
        // two-hidden-layer items keep the accumulator: layer 2 accumulates into it (see below)
        named_bar_arrive(BAR_HID, kEpiThreads + 32);
        if (T.n_hidden != 2) warp_arrive(&bars->acc1_empty[buf]);
        if (warp == 0) RFK_TRACE(3, 2);
        // ---- deferred softmax of the previous item, overlapping this item's layer-2 MMAs
        if (have_prev) finish_prev();
        // ---- layer 2 (two-hidden-layer trials): acc2 -> hidden
        if (T.n_hidden == 2) {
          wait_prof<kProf>(&bars->acc2_full, acc2_phase, prof, PF_EPI_ACC2);
          if (warp == 0) RFK_TRACE(3, 3);
          acc2_phase ^= 1u;
          tc_fence_after_sync();
          const int split = T.h2_pad < 64 ? T.h2_pad : 64;
          // layer 2 accumulated into this item's (drained) layer-1 buffer
          epilogue_hidden<kF16>(lane_addr + COL_ACC1 + buf * HMAX, half ? split : 0, half ? T.h2_pad : split, bias + HMAX,
                                1.0f, hid_row_addr);
          tc_fence_before_sync();
          named_bar_arrive(BAR_HID, kEpiThreads + 32);
          warp_arrive(&bars->acc1_empty[buf]);
          if (warp == 0) RFK_TRACE(3, 4);
        }
        have_prev = true;
        prev_tile = tile;
        prev_t = t;
        prev_buf = buf;
      }
    }
    if (have_prev) {
      fetch_prev_logits((uint32_t)((item - 1) & 1));
      tc_fence_before_sync();
      finish_prev();
    }
  } else if (warp < kEpiWarps + kConvWarps) {
    // =============================================================== converter warps
    // One step = 128 K elements of one item: 128 rows x 128 bytes of uint8 in the raw ring (written by
    // TMA with the 128-byte swizzle: 16-byte chunk c of row r sits at chunk position c ^ (r & 7), so
    // the eight lanes of a quarter-warp, which read the same chunk of eight consecutive rows, hit
    // eight different bank groups).  A thread owns half a row (warp % 4 = the TMEM lane quarter of
    // the row, warp / 4 = which 64 bytes): 4 x LDS.128, 32 exact conversions, and the 64 operand
    // values go straight into TENSOR memory (tcgen05.st), where the layer-1 MMAs read their A
    // operand -- the converted tile never touches shared memory.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsConv));
    const int quarter = warp & 3;
    const int khalf = (warp - kEpiWarps) >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t raw_row = smem_base + OFF_RAW + (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    uint32_t raw_chunk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) raw_chunk[j] = raw_row + ((((uint32_t)(4 * khalf + j)) ^ sw) << 4);
    const uint32_t a_dst0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + COL_A + (uint32_t)khalf * 32u;
    int rstage = 0;
    uint32_t rphase = 0;
    int gs = 0;                                    // global step counter of this CTA
    for (int tile = tile0; tile < P.n_tiles; tile += tile_step) {
      for (int t = t0; t < t1; ++t) {
        for (int kb = 0; kb < n_kb; ++kb, ++gs) {
          // 16-byte chunks of this thread's half row that the MMAs will read (the row may end inside the step)
          int live = (P.d - kb * KB - khalf * 64 + 15) >> 4;
          live = (P.debug & 2) ? 0 : (live < 0 ? 0 : (live > 4 ? 4 : live));
          // (1) raw bytes -> operand pairs in registers, as soon as the raw tile has landed: the A stage
          //     may still be in use by the MMAs of step gs - A_STAGES, so nothing is stored yet
          mbar_wait(&bars->raw_full[rstage], rphase);
          if (warp == kEpiWarps) RFK_TRACE(0, 1);
          const uint32_t rsrc = (uint32_t)rstage * RAW_STAGE;
          uint32_t p[32];
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            if (live > 2 * h2) {                   // warp-uniform
              const uint4 r0 = lds_u4(raw_chunk[2 * h2] + rsrc), r1 = lds_u4(raw_chunk[2 * h2 + 1] + rsrc);
              uint32_t* q = p + 16 * h2;
              q[0] = u8x2_to_op<kF16>(r0.x, 0);  q[1] = u8x2_to_op<kF16>(r0.x, 2);
              q[2] = u8x2_to_op<kF16>(r0.y, 0);  q[3] = u8x2_to_op<kF16>(r0.y, 2);
              q[4] = u8x2_to_op<kF16>(r0.z, 0);  q[5] = u8x2_to_op<kF16>(r0.z, 2);
              q[6] = u8x2_to_op<kF16>(r0.w, 0);  q[7] = u8x2_to_op<kF16>(r0.w, 2);
              q[8] = u8x2_to_op<kF16>(r1.x, 0);  q[9] = u8x2_to_op<kF16>(r1.x, 2);
              q[10] = u8x2_to_op<kF16>(r1.y, 0); q[11] = u8x2_to_op<kF16>(r1.y, 2);
              q[12] = u8x2_to_op<kF16>(r1.z, 0); q[13] = u8x2_to_op<kF16>(r1.z, 2);
              q[14] = u8x2_to_op<kF16>(r1.w, 0); q[15] = u8x2_to_op<kF16>(r1.w, 2);
            }
          }
          // the raw slot is free as soon as its bytes sit in registers (the conversions above consumed the loads)
          warp_arrive(&bars->raw_empty[rstage]);
          // (2) A stage gs % A_STAGES was last read by the MMAs of step gs - A_STAGES
          if (gs >= A_STAGES) mbar_wait(&bars->step_done[(gs - A_STAGES) % B_STAGES], (uint32_t)(((gs - A_STAGES) / B_STAGES) & 1));
          if (warp == kEpiWarps) RFK_TRACE(0, 2);
          tc_fence_after_sync();
          const uint32_t dst = a_dst0 + (uint32_t)(gs % A_STAGES) * (KB / 2);
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2)
            if (live > 2 * h2) tmem_st16(dst + 16u * h2, *reinterpret_cast<const uint32_t(*)[16]>(p + 16 * h2));
          tmem_st_wait();
          tc_fence_before_sync();
          // (3) hand the stage to the issuer: a hardware named barrier (the issuer does bar.sync on the
          //     same id) costs a few tens of cycles, an mbarrier arrive + waiter wake-up several hundred
          named_bar_arrive(BAR_A0 + gs % A_STAGES, kConvThreads + 32);
          if (warp == kEpiWarps) RFK_TRACE(0, 3);
          if (++rstage == RAW_STAGES) { rstage = 0; rphase ^= 1u; }
        }
      }
    }
  } else {
    // warpgroup 4: loader, layer-1 MMA issuer, spare, layers-2/3 MMA issuer -- all give registers back
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsLight));
  }
  if (warp == WARP_LOAD) {
    // =============================================================== loader: weights, raw query tiles, parameters
    // Per step j, in the order the two slots come free: the raw uint8 tile of step j (TMA tensor load
    // into the RAW_STAGES-deep raw ring, as soon as the conversion of step j - 3 has read its slot)
    // and the W1 block of step j (cp.async.bulk of the pre-swizzled block into the B_STAGES-deep
    // weight ring, as soon as the MMAs of step j - 4 have completed).
    // Per-item parameters (biases, W3: double buffered by item parity; W2: one buffer, released
    // as soon as the layer-2 MMAs of the previous item have completed) ride along as a third,
    // non-blocking cursor that is polled once per step; with a single trial nothing is reloaded.
    int my_tiles = 0;
    for (int tile = tile0; tile < P.n_tiles; tile += tile_step) ++my_tiles;
    const int total_items = (P.debug & 64) ? 0 : my_tiles * w;
    const int n_atoms = (P.d + ATOM_K - 1) / ATOM_K;
    int p_item = 0, p_t = t0, p_sub = 0, n2 = 0;
    auto param_poll = [&](bool blocking) {
      // at most one action per call; returns when the cursor cannot advance (non-blocking mode)
      if (p_item >= total_items) return;
      const TrialDev& T = P.trial[p_t];
      const int buf = p_item & 1;
      const uint32_t par_free2 = (uint32_t)(((p_item >> 1) & 1) ^ 1);
      if (p_sub < 2 && !(w > 1 || p_item < 2)) p_sub = 2;
      if (p_sub == 0) {
        if (blocking) mbar_wait(&bars->bias_free[buf], par_free2);
        else if (!mbar_test(&bars->bias_free[buf], par_free2)) return;
        bulk_g2s_expect_elect(smem + OFF_BIAS + buf * BIAS_BYTES, T.bias, BIAS_BYTES, &bars->bias_full[buf]);
        p_sub = 1;
        return;
      }
      if (p_sub == 1) {
        if (blocking) mbar_wait(&bars->w3_free[buf], par_free2);
        else if (!mbar_test(&bars->w3_free[buf], par_free2)) return;
        const int hl_pad = T.n_hidden == 2 ? T.h2_pad : T.h1_pad;
        const uint32_t w3_bytes = (uint32_t)((hl_pad + 63) / 64) * CPAD * 128u;
        bulk_g2s_expect_elect(smem + OFF_W3 + buf * W3_BYTES, T.w3, w3_bytes, &bars->w3_full[buf]);
        p_sub = 2;
        return;
      }
      if (T.n_hidden == 2 && (w > 1 || n2 == 0)) {
        if (blocking) mbar_wait(&bars->w2_free, (uint32_t)((n2 & 1) ^ 1));
        else if (!mbar_test(&bars->w2_free, (uint32_t)((n2 & 1) ^ 1))) return;
        const uint32_t w2_bytes = (uint32_t)((T.h1_pad + 63) / 64) * (uint32_t)T.h2_pad * 128u;
        bulk_g2s_expect_elect(smem + OFF_W2, T.w2, w2_bytes, &bars->w2_full);
      }
      if (T.n_hidden == 2) ++n2;
      ++p_item;
      p_sub = 0;
      if (++p_t == t1) p_t = t0;
    };
    // Waiting for a ring slot must not starve the parameter cursor (the MMAs the slot waits for may
    // themselves be waiting, through the epilogue, for parameters): poll it between bounded waits.
    auto wait_serving = [&](uint64_t* bar, uint32_t parity) {
      if (mbar_try_wait(bar, parity)) return;
      uint32_t spins = 0;
      while (!mbar_try_wait_hint(bar, parity, 200u)) {
        param_poll(false);
        if (++spins > (1u << 24)) {
          if (lane == 0) printf("rafiki_b200: loader wait timed out (block %d)\n", blockIdx.x);
          __trap();
        }
      }
    };
    // the parameters of the first two items have free buffers: load them before anything else
    for (int i = 0; i < 6; ++i) param_poll(false);
    int bstage = 0, rstage = 0;
    uint32_t bphase = 0, rphase = 0;
    for (int tile = tile0; tile < P.n_tiles; tile += tile_step) {
      for (int t = t0; t < t1; ++t) {
        const TrialDev& T = P.trial[t];
        const uint32_t atom_bytes = (uint32_t)T.h1_pad * 128u;        // one 64-wide K atom of W1
        const uint8_t* src = T.w1;
        for (int kb = 0; kb < n_kb; ++kb) {
          wait_serving(&bars->raw_empty[rstage], rphase ^ 1u);
          RFK_TRACE(1, 1);
          if (P.debug & 32)
            mbar_arrive_elect(&bars->raw_full[rstage]);
          else
            tma_load_2d_expect_elect(smem + OFF_RAW + rstage * RAW_STAGE, &x_map, kb * KB, tile * TILE_M, RAW_STAGE,
                                     &bars->raw_full[rstage]);
          const uint32_t bytes = atom_bytes * (uint32_t)(n_atoms - 2 * kb < 2 ? n_atoms - 2 * kb : 2);
          wait_serving(&bars->step_done[bstage], bphase ^ 1u);
          RFK_TRACE(1, 3);
          if (P.debug & 1)
            mbar_arrive_elect(&bars->b_full[bstage]);
          else
            bulk_g2s_expect_elect(smem + OFF_B + bstage * B_STAGE, src + (size_t)kb * 2 * atom_bytes, bytes, &bars->b_full[bstage]);
          if (++bstage == B_STAGES) { bstage = 0; bphase ^= 1u; }
          if (++rstage == RAW_STAGES) { rstage = 0; rphase ^= 1u; }
          param_poll(false);
        }
      }
    }
    while (p_item < total_items) param_poll(true);
  } else if (warp == WARP_MMA) {
    // =============================================================== MMA issuer, layer 1
    // Executed by all 32 lanes in lock step (mbarrier waits included); the *_elect helpers pick
    // one lane per statement, so there is no divergent region.  A comes from tensor memory (the
    // converters' ring, 8 columns per 16-element K step), B from shared memory: the descriptor is a
    // constant upper word plus (address >> 4) in the lower word.  One commit per step releases both
    // the weight stage (to the loader) and the A stage (to the converters).
    const uint32_t b_lo0 = desc_lo_sw128(smem_base + OFF_B);
    int item = 0, gs = 0;
    for (int tile = tile0; tile < P.n_tiles; tile += tile_step) {
      for (int t = t0; t < t1; ++t, ++item) {
        const TrialDev& T = P.trial[t];
        const int buf = item & 1;
        const uint32_t idesc1 = idesc_16bit_f32(kF16, TILE_M, T.h1_pad);
        const uint32_t d_tmem = tmem_base + COL_ACC1 + buf * HMAX;
        wait_prof<kProf>(&bars->acc1_empty[buf], (uint32_t)(((item >> 1) & 1) ^ 1), prof, PF_MMA_ACC1_EMPTY);
        for (int kb = 0; kb < n_kb; ++kb, ++gs) {
          const int as = gs % A_STAGES, bs = gs % B_STAGES;
          wait_prof<kProf>(&bars->b_full[bs], (uint32_t)((gs / B_STAGES) & 1), prof, PF_MMA_W23_FULL);
          {
            const long long tw = kProf ? clock64() : 0;
            named_bar_sync(BAR_A0 + as, kConvThreads + 32);            // the converters have filled A stage `as`
            if (kProf) prof[PF_MMA_FULL] += clock64() - tw;
          }
          RFK_TRACE(2, 1);
          tc_fence_after_sync();
          if (!(P.debug & 4)) {
            // two 64-wide atoms per step; the ragged end of the row may leave the second one short or empty
            const int k16 = (P.d - kb * KB) >> 4;                      // 16-element K steps left in this row
            const uint32_t a = tmem_base + COL_A + (uint32_t)as * (KB / 2);
            const uint32_t b = b_lo0 + (uint32_t)bs * (B_STAGE >> 4);
            mma4_f16_ts_elect(d_tmem, a, desc_from_lo(b), idesc1, kb != 0, k16 < 4 ? k16 : 4);
            if (k16 > 4)
              mma4_f16_ts_elect(d_tmem, a + 32u, desc_from_lo(b + (((uint32_t)T.h1_pad * 128u) >> 4)), idesc1, 1u,
                                k16 - 4 < 4 ? k16 - 4 : 4);
          }
          RFK_TRACE(2, 2);
          mma_commit_elect(&bars->step_done[bs]);
          RFK_TRACE(2, 3);
        }
        mma_commit_elect(&bars->acc1_full[buf]);
      }
    }
  } else if (warp == WARP_TAIL) {
    // =============================================================== MMA issuer, layers 2 and 3
    // hidden tile (A, written by the epilogue warps) x W2 / W3 (B, parameter loader).  A separate
    // warp, so that the layer-1 issuer never waits for an epilogue; the tensor pipe interleaves
    // the two instruction streams, which touch disjoint accumulators.
    const uint32_t w2_lo = desc_lo_sw128(smem_base + OFF_W2), w3_lo = desc_lo_sw128(smem_base + OFF_W3);
    auto issue_small_layer = [&](uint32_t d_tmem, uint32_t b_lo, uint32_t b_atom, uint32_t idesc, int ksteps) {
      // A = hidden activations in tensor memory (8 columns per K step); B in 64-wide atoms of 4 steps
      for (int k0 = 0; k0 < ksteps; k0 += 4) {
        const uint32_t a = tmem_base + COL_HID + (uint32_t)k0 * 8u;
        const uint32_t b = b_lo + (uint32_t)(k0 >> 2) * b_atom;
        mma4_f16_ts_elect(d_tmem, a, desc_from_lo(b), idesc, k0 != 0, ksteps - k0 < 4 ? ksteps - k0 : 4);
      }
    };
    int item = 0, n2 = 0;
    for (int tile = tile0; tile < ((P.debug & 64) ? 0 : P.n_tiles); tile += tile_step) {
      for (int t = t0; t < t1; ++t, ++item) {
        const TrialDev& T = P.trial[t];
        const int buf = item & 1;
        if (T.n_hidden == 2) {
          if (w > 1 || n2 == 0) wait_prof<kProf>(&bars->w2_full, (uint32_t)(n2 & 1), prof, PF_MMA_W23_FULL);
          ++n2;
          {
            const long long tw = kProf ? clock64() : 0;
            named_bar_sync(BAR_HID, kEpiThreads + 32);
            if (kProf) prof[PF_MMA_HID_L2] += clock64() - tw;
          }
          RFK_TRACE(4, 1);
          tc_fence_after_sync();
          issue_small_layer(tmem_base + COL_ACC1 + (uint32_t)buf * HMAX, w2_lo, ((uint32_t)T.h2_pad * 128u) >> 4,
                            idesc_16bit_f32(kF16, TILE_M, T.h2_pad), T.h1_pad >> 4);
          mma_commit_elect(&bars->acc2_full);
          mma_commit_elect(&bars->w2_free);
          RFK_TRACE(4, 2);
        }
        if (w > 1 || item < 2) wait_prof<kProf>(&bars->w3_full[buf], (uint32_t)((item >> 1) & 1), prof, PF_MMA_W23_FULL);
        {
          const long long tw = kProf ? clock64() : 0;
          named_bar_sync(BAR_HID, kEpiThreads + 32);
          if (kProf) prof[PF_MMA_HID_L3] += clock64() - tw;
        }
        RFK_TRACE(4, 3);
        tc_fence_after_sync();
        const int hl_pad = T.n_hidden == 2 ? T.h2_pad : T.h1_pad;
        issue_small_layer(tmem_base + COL_ACC3, w3_lo + (uint32_t)buf * (W3_BYTES >> 4), (CPAD * 128u) >> 4,
                          idesc_16bit_f32(kF16, TILE_M, CPAD), hl_pad >> 4);
        mma_commit_elect(&bars->acc3_full);
        mma_commit_elect(&bars->w3_free[buf]);
        RFK_TRACE(4, 4);
      }
    }
  }
  if (kProf && P.prof != nullptr && lane == 0 &&
      (warp == 0 || warp == kEpiWarps || warp == WARP_LOAD || warp == WARP_MMA || warp == WARP_TAIL)) {
    const long long total = clock64() - t_start;
    if (warp == 0) prof[PF_EPI_TOTAL] = total;
    if (warp == kEpiWarps) prof[PF_CONV_TOTAL] = total;
    if (warp == WARP_MMA) prof[PF_MMA_TOTAL] = total;
    for (int i = 0; i < PF_COUNT; ++i)
      if (prof[i] != 0) atomicAdd(reinterpret_cast<unsigned long long*>(P.prof + i), (unsigned long long)prof[i]);
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
  if (h->dtype != RFK_BF16 && h->dtype != RFK_F16) return false;
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
  pack_kmajor_sw128(l1.w16_host.data(), l1.out_pad, l1.in_pad, &img);
  int rc = upload_vec(&s->w1, img);
  if (rc != RFK_OK) return rc;
  std::vector<float> bias(BIAS_FLOATS, 0.f);
  for (int j = 0; j < l1.out_pad; ++j) bias[j] = l1.b_fold_host[j];
  if (h->n_hidden == 2) {
    const Layer& l2 = h->layers[1];
    pack_kmajor_sw128(l2.w16_host.data(), l2.out_pad, l2.in_pad, &img);
    rc = upload_vec(&s->w2, img);
    if (rc != RFK_OK) return rc;
    for (int j = 0; j < l2.out_pad; ++j) bias[HMAX + j] = l2.b_fold_host[j];
  }
  pack_kmajor_sw128(l3.w16_host.data(), l3.out_pad, l3.in_pad, &img);
  rc = upload_vec(&s->w3, img);
  if (rc != RFK_OK) return rc;
  for (int j = 0; j < l3.out_pad && j < CPAD; ++j) bias[2 * HMAX + j] = l3.b_fold_host[j];
  for (int j = h->n_classes; j < CPAD; ++j) bias[2 * HMAX + j] = -INFINITY;     // padded classes: see softmax16
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
  {
    // the setmaxnreg budget above assumes the registers the compiler really gave the kernel
    cudaFuncAttributes fa;
    RFK_CUDA_TRY(cudaFuncGetAttributes(&fa, fused_ensemble_kernel<true, false>));
    if (fa.numRegs * (kEpiWarps + kConvWarps + 4) < kEpiWarps * kRegsEpi + kConvWarps * kRegsConv + 4 * kRegsLight) {
      set_error("fused kernel was compiled with %d registers per thread; its setmaxnreg budget needs %d", fa.numRegs,
                kRegsLaunch);
      return RFK_ERR_UNSUPPORTED;
    }
  }
  RFK_REQUIRE(h->device >= 0 && h->device < 64, "fused path: device ordinal %d not supported", h->device);
  static std::atomic<bool> attr_set[64];
  if (!attr_set[h->device].load(std::memory_order_acquire)) {     // setting it twice is harmless

    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_ensemble_kernel<true, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_ensemble_kernel<false, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_ensemble_kernel<true, true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_set[h->device].store(true, std::memory_order_release);
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

bool fused_images(const rfk_mlp* h, FusedImages* out) {
  const FusedState* s = static_cast<const FusedState*>(h->fused);
  if (s == nullptr) return false;
  out->w1 = s->dev.w1;
  out->w2 = s->dev.w2;
  out->w3 = s->dev.w3;
  out->bias = s->dev.bias;
  out->h1_pad = s->dev.h1_pad;
  out->h2_pad = s->dev.h2_pad;
  out->n_hidden = s->dev.n_hidden;
  return true;
}

size_t fused_workspace_bytes(int w, int n_classes, int64_t n, int d) {
  // large batches: the pair kernel writes per-trial probabilities for kernel 3
  if (pair_applicable(d, n)) return pair_workspace_bytes(w, n_classes, n);
  // split mode (see the kernel): used when all (tile, trial) items fit on the GPU at once
  const int64_t n_tiles = (n + TILE_M - 1) / TILE_M;
  if (w > 1 && n_tiles * w <= (int64_t)num_sms()) return (size_t)w * (size_t)n * n_classes * sizeof(float);
  return 0;
}

int fused_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n, float* out,
                  void* ws, size_t ws_bytes, cudaStream_t st) {
  RFK_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "fused forward: x must be 16-byte aligned");
  const int d = trials[0]->d;
  // large batches go to the CTA-pair kernel (weights resident in shared memory); an ensemble needs the
  // per-trial scratch for that, without it the single-CTA kernel below still answers
  if (pair_applicable(d, n) &&
      (w == 1 || (ws != nullptr && ws_bytes >= pair_workspace_bytes(w, trials[0]->n_classes, n))))
    return pair_forward(trials, w, total_trials, x, n, out, ws, ws_bytes, st);
  // TMA descriptor of the query batch: uint8 [n, d]; one box = 128 rows x 128 bytes, 128-byte swizzle
  CUtensorMap x_map;
  {
    const int rc = make_tmap_u8_2d(&x_map, x, (uint64_t)n, (uint64_t)d, (uint64_t)d, TILE_M, KB, true);
    if (rc != RFK_OK) return rc;
  }
  FusedParams P;
  memset(&P, 0, sizeof(P));
  for (int t = 0; t < w; ++t) {
    const FusedState* s = static_cast<const FusedState*>(trials[t]->fused);
    RFK_REQUIRE(s != nullptr, "fused forward: trial %d was not packed for the fused path", t);
    P.trial[t] = s->dev;
  }
  P.out = out;
  P.n = n;
  P.d = trials[0]->d;
  P.n_kb = (P.d + KB - 1) / KB;
  P.w = w;
  P.total_trials = total_trials;
  P.n_classes = trials[0]->n_classes;
  P.n_tiles = (int)((n + TILE_M - 1) / TILE_M);
  {
    const char* dbg = getenv("RFK_FUSED_DEBUG");
    P.debug = dbg ? atoi(dbg) : 0;
  }
  const int sms = num_sms();
  int grid = P.n_tiles < sms ? P.n_tiles : sms;
  const size_t split_bytes = w > 1 && (int64_t)P.n_tiles * w <= (int64_t)sms ? (size_t)w * (size_t)n * P.n_classes * sizeof(float) : 0;
  const bool split = split_bytes != 0 && ws != nullptr && ws_bytes >= split_bytes && getenv("RFK_NO_SPLIT") == nullptr;
  if (split) {
    P.split = 1;
    P.out = static_cast<float*>(ws);
    grid = P.n_tiles * w;
  }
  const char* prof_env = getenv("RFK_FUSED_PROF");
  if (prof_env != nullptr && atoi(prof_env) != 0 && trials[0]->dtype == RFK_F16) {
    // development aid: synchronous, allocates; prints the per-role blocked-cycle breakdown
    long long* dprof = nullptr;
    const size_t prof_elems = PF_COUNT + 5 * kTraceLen;
    RFK_CUDA_TRY(cudaMalloc(&dprof, prof_elems * sizeof(long long)));
    RFK_CUDA_TRY(cudaMemsetAsync(dprof, 0, prof_elems * sizeof(long long), st));
    P.prof = dprof;
    P.trace = atoi(prof_env) >= 2 ? dprof + PF_COUNT : nullptr;
    fused_ensemble_kernel<true, true><<<grid, kThreads, SMEM_BYTES, st>>>(P, x_map);
    std::vector<long long> hall(prof_elems);
    RFK_CUDA_TRY(cudaMemcpyAsync(hall.data(), dprof, prof_elems * sizeof(long long), cudaMemcpyDeviceToHost, st));
    RFK_CUDA_TRY(cudaStreamSynchronize(st));
    cudaFree(dprof);
    const long long* hprof = hall.data();
    if (P.trace != nullptr) {
      const char* path = getenv("RFK_FUSED_TRACE_FILE");
      FILE* f = fopen(path ? path : "fused_trace.txt", "w");
      if (f) {
        static const char* roles[5] = {"conv", "load", "mma", "epi", "tail"};
        for (int r = 0; r < 5; ++r)
          for (int i = 0; i < kTraceLen; ++i) {
            const long long v = hall[PF_COUNT + r * kTraceLen + i];
            if (v != 0) fprintf(f, "%s %d %lld %lld\n", roles[r], i, v >> 4, v & 15);
          }
        fclose(f);
      }
    }
    static const char* names[PF_COUNT] = {"epi:bias_full", "epi:acc1_full", "epi:acc2_full", "epi:acc3_full",
                                          "epi:TOTAL", "conv:raw_full", "conv:empty", "conv:TOTAL", "load:empty",
                                          "load:raw_empty", "par:bias_free", "par:w23_free", "mma:acc1_empty",
                                          "mma:full", "mma:w23_full", "mma:hid_ready(L2)", "mma:hid_ready(L3)",
                                          "mma:TOTAL"};
    const double items = (double)P.n_tiles * w;
    fprintf(stderr, "[rfk prof] grid %d, items %g; cycles per item (avg over CTAs):\n", grid, items);
    for (int i = 0; i < PF_COUNT; ++i) fprintf(stderr, "[rfk prof]   %-18s %10.0f\n", names[i], hprof[i] / items);
    return RFK_OK;
  }
  {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = getenv("RFK_NO_PDL") == nullptr ? 1 : 0;
    if (trials[0]->dtype == RFK_F16)
      RFK_CUDA_TRY(cudaLaunchKernelEx(&cfg, fused_ensemble_kernel<true, false>, P, x_map));
    else
      RFK_CUDA_TRY(cudaLaunchKernelEx(&cfg, fused_ensemble_kernel<false, false>, P, x_map));
  }
  RFK_LAUNCH_CHECK("fused_ensemble_kernel");
  if (split) {
    // kernel 3 over the W per-trial outputs: same float64 sum and rounding as the in-kernel mean
    const float* ptrs[RFK_MAX_TRIALS];
    for (int t = 0; t < w; ++t) ptrs[t] = static_cast<const float*>(ws) + (size_t)t * (size_t)n * P.n_classes;
    return ensemble_partial_mean(ptrs, w, n, P.n_classes, out, (double)total_trials, st);
  }
  return RFK_OK;
}

}  // namespace rfk
