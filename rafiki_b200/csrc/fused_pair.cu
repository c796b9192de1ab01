// Kernels 1+2 for large batches on CTA PAIRS: the trial's weights stay in shared memory.
//
// Same arithmetic and the same reference code replaced as fused_trial.cu (InferenceWorker predict,
// rafiki/worker/inference.py:53 -> TfFeedForward.predict, TfFeedForward.py:89-97,141-157); what
// changes is where the bytes travel.  fused_trial.cu streams the first-layer weights (200 KB per
// 128-query tile and trial) from L2 into every SM again for every tile: at batch 65536 x 8 trials
// that is 1.2 GB through L2 per launch, and the kernel runs at the chip's L2 -> SM throughput
// (about 42 B/clk per SM measured: 300 KB per item / 7100 cycles), not at the tensor pipe's.
// Here two CTAs of a cluster work as one 256-query tile (tcgen05 cta_group::2, M = 256): each
// CTA keeps HALF of the hidden units' weight rows (W1: 104 KB for d 784, W2: 16 KB, W3: 2 KB) in
// shared memory for as long as the pair stays on one trial, and the only per-tile traffic is the
// uint8 tile itself.  The loop order is therefore trial-outer: a pair takes a contiguous range of
// (trial, 256-query tile) items in trial-major order, reloads the weights when the trial changes
// (at most once or twice per launch), and writes per-trial probabilities; the cross-trial mean is
// kernel 3 (ensemble.cu) over those, in float64 in trial order with one rounding -- the same sum
// the single-CTA kernel keeps in registers.
//
// Roles per CTA (640 threads, as in fused_trial.cu): epilogue x 8 warps (hidden layers; the half-0
// warps also turn logits into staged probabilities, the half-1 warps store the staged tile), converters
// x 8 (uint8 -> fp16 into tensor memory; a warp converts its rows of a whole raw stage and hands its
// two K atoms over one at a time, two groups of four warps alternate stages), loader (TMA of raw
// tiles; bulk copies of the weights at a trial change), layer-1 MMA issuer and layers-2/3 MMA issuer
// (LEADER CTA only: one thread issues for both CTAs), and a warp that forwards "weights landed" to the
// leader and, in the multi-GPU step, carries the previous batch's cross-rank sum.  Hand-offs that
// cross the pair (operands ready, accumulator drained) are arrivals on the leader's mbarriers through
// the cluster window; completions come back as tcgen05.commit multicast to both CTAs.  No load here
// is a multicast load: every CTA fetches its own bytes.  The kernel is launched as a programmatic
// dependent of the previous kernel of its stream (griddep_wait below).
#include <stdlib.h>

#include <atomic>
#include <type_traits>

#include "fused_common.cuh"
#include "pack.cuh"
#include "rfk_internal.cuh"
#include "sm100.cuh"
#include "tmap.cuh"

namespace {

using namespace sm100;
using namespace fusedk;

constexpr int TILE_M = 128;          // rows per CTA; a pair covers 256
constexpr int ATOM_K = 64;
constexpr int KB = 128;              // K elements per raw stage (one TMA box)
#ifndef RFK_DEV_A_STAGES
#define RFK_DEV_A_STAGES 5
#endif
#ifndef RFK_DEV_ACC_STRIDE          // development: 64 frees tensor memory for a longer A ring (valid for <= 64 hidden units only)
#define RFK_DEV_ACC_STRIDE 128
#endif
constexpr int A_STAGES = RFK_DEV_A_STAGES;   // A operand ring in tensor memory: stages of ONE 64-wide K atom
constexpr int A_COLS = ATOM_K / 2;   // 32 columns per stage
constexpr int MAX_RAW_STAGES = 6;
constexpr int MIN_RAW_STAGES = 3;
constexpr int HMAX = 128;
constexpr int HHALF = HMAX / 2;      // weight rows (hidden units) per CTA
constexpr int CPAD = 16;
constexpr int W1_ATOM = HHALF * 128;              // 8 KB: one 64-wide K atom of this CTA's half of W1
constexpr int W2_BYTES = 2 * HHALF * 128;         // 16 KB
constexpr int W3_ATOM = (CPAD / 2) * 128;         // 1 KB: 8 class rows per CTA
constexpr int W3_BYTES = 2 * W3_ATOM;
constexpr int RAW_STAGE = TILE_M * KB;            // 16 KB of uint8
constexpr int BIAS_FLOATS = HMAX + HMAX + CPAD;   // b1 | b2 | b3
constexpr int BIAS_BYTES = BIAS_FLOATS * 4;
constexpr int OUT_BYTES = TILE_M * CPAD * 4;
constexpr int SMEM_MAX = 227 * 1024;

constexpr int kEpiWarps = 8, kConvWarps = 8;
constexpr int kThreads = (kEpiWarps + kConvWarps + 4) * 32;   // 640
// role indices (see the mapping from physical warps in the kernel)
constexpr int WARP_LOAD = kEpiWarps + kConvWarps, WARP_MMA = WARP_LOAD + 1, WARP_FWD = WARP_LOAD + 2,
              WARP_TAIL = WARP_LOAD + 3;
constexpr int kRegsLaunch = 96;
constexpr int kRegsEpi = 136, kRegsConv = 72, kRegsLight = 64;
// setmaxnreg.inc only hands out what the CTA's own warps have given back (the launch allocation)
static_assert(kEpiWarps * kRegsEpi + kConvWarps * kRegsConv + 4 * kRegsLight <= (kEpiWarps + kConvWarps + 4) * kRegsLaunch,
              "setmaxnreg budget exceeds the launch allocation");
constexpr int BAR_STAGE_FULL = 1, BAR_STAGE_FREE = 2;

constexpr int TMEM_COLS = 512;
constexpr int COL_ACC1 = 0;     // 2 x 128: layer-1 accumulators (layer 2 accumulates into the drained buffer)
constexpr int ACC_STRIDE = RFK_DEV_ACC_STRIDE;
constexpr int COL_A = 2 * ACC_STRIDE;      // A ring: A_STAGES x 32 columns
#ifdef RFK_DEV_TIMING_ALIAS_ACC3     // development, TIMING ONLY (results are wrong): logits land in the layer-1 accumulator
constexpr int COL_HID = 448;
constexpr int COL_ACC3 = 0;
static_assert(COL_A + A_STAGES * A_COLS <= COL_HID && COL_HID + HMAX / 2 <= TMEM_COLS, "tensor memory map");
#else
constexpr int COL_HID = 416;    // hidden activations (16-bit pairs): 64 columns
constexpr int COL_ACC3 = 480;
static_assert(COL_A + A_STAGES * A_COLS <= COL_HID && COL_HID + HMAX / 2 <= COL_ACC3 && COL_ACC3 + CPAD <= TMEM_COLS,
              "tensor memory map");
#endif

struct PairTrial {
  const uint8_t* w1;    // [n_atoms][h1_pad rows][128 B] swizzled (the single-CTA kernel's image)
  const uint8_t* w2;    // [h1_pad / 64][h2_pad rows][128 B]
  const uint8_t* w3;    // [hl_pad / 64][16 rows][128 B]
  const float* bias;    // [BIAS_FLOATS]
  int h1_pad, h2_pad, n_hidden, pad_;
};

struct PairParams {
  PairTrial trial[RFK_MAX_TRIALS];
  float* out;                 // trial t writes out + t * trial_stride
  long long trial_stride;     // elements between the outputs of consecutive trials (0: a single trial)
  long long n;
  double divisor;             // the stored value is float(p / divisor)
  float div_f, div_recip;     // divisor as float, and 1 / divisor when it is a power of two
  int div_mode;               // 0: divisor 1, 1: a power of two (multiply by div_recip), 2: one correctly rounded division
  int d, n_kb, n_atoms, w, n_classes, n_pt, raw_stages;
  int items_q, items_r;       // items per pair: pair p takes items_q (+ 1 if p < items_r), in order
  int off_w2, off_w3, off_raw, off_bias, off_out, off_bar;
  rfk::PeerHook hook;   // all zero: no hand-shake (see rfk_internal.cuh)
  int debug;            // development (RFK_PAIR_DEBUG, timing experiments only): 4 skip the layer-1 MMAs, 8 no raw ring,
                        // 16 no tensor-memory stores of A, 32 no softmax / output, 64 no conversion
  unsigned int* prof;   // development (RFK_PAIR_PROF): per-warp segment cycle sums, [2 CTAs][20 warps][12]
  long long* trace;     // development (RFK_PAIR_TRACE): event timestamps of pair 0, [2 CTAs][5 roles][kTraceLen]
};
constexpr int kTraceLen = 1024;
#define PTRACE(role, tag)                                                                              \
  do {                                                                                               \
    if (kTrace && P.trace != nullptr && blockIdx.x < 2 && lane == 0 && trace_n < kTraceLen) {          \
      P.trace[((int)blockIdx.x * 5 + (role)) * kTraceLen + trace_n] = (clock64() << 4) | (long long)(tag); \
      ++trace_n;                                                                                     \
    }                                                                                                \
  } while (0)

// Development profile (kDev, RFK_PAIR_PROF): cycles per segment of a role's loop, summed in registers
// (no stores inside the loops), written once per warp at the end: prof[cta][warp][8].
#define PROF(k)                                   \
  do {                                            \
    if (kDev) {                                   \
      const uint32_t now_ = (uint32_t)clock();    \
      prof_acc[k] += now_ - prof_t;               \
      prof_t = now_;                              \
    }                                             \
  } while (0)

struct PBars {
  uint64_t raw_full[MAX_RAW_STAGES], raw_empty[MAX_RAW_STAGES];   // own CTA
  uint64_t a_full[A_STAGES];      // leader: 2 x 4 converter warps (the group that owns the atom's 64 bytes of every row)
  uint64_t a_free[A_STAGES];      // both: the MMAs that read the stage are complete (multicast commit)
  uint64_t acc1_full[2];          // both (multicast commit)
  uint64_t acc1_empty[2];         // leader: 2 x 8 epilogue warps
  uint64_t hid_ready;             // leader: 2 x 8 epilogue warps
  uint64_t acc2_full, acc3_full;  // both (multicast commit)
  uint64_t w_full;                // own CTA: this trial's weights landed (transaction bytes)
  uint64_t w_free;                // own CTA: 8 epilogue warps are done with the previous trial's weights
  uint64_t w_ready;               // leader: both CTAs' weights landed
};
constexpr int TAIL_BYTES = 10 * 1024;   // bias + output staging + barriers + TMEM slot, after the raw ring
static_assert(((BIAS_BYTES + 15) / 16 * 16) + OUT_BYTES + (int)sizeof(PBars) + 16 <= TAIL_BYTES, "tail region");

// The items of a pair in order, without a division in the kernel (a 64-bit division is a call, and
// values that come back from a call are no longer warp-uniform for the compiler: the MMA issue then
// falls off the uniform datapath).
struct ItemIter {
  int t, pt;
  __device__ __forceinline__ void init(int first, int n_pt) {
    t = 0;
    pt = first;
    while (pt >= n_pt) {
      pt -= n_pt;
      ++t;
    }
  }
  __device__ __forceinline__ void next(int n_pt) {
    if (++pt == n_pt) {
      pt = 0;
      ++t;
    }
  }
};

// leader-side arrival through the cluster window: one per warp (see warp_arrive in fused_common.cuh)
__device__ __forceinline__ void warp_arrive_leader(uint32_t leader_bar_addr) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive_cluster(leader_bar_addr);
}

// plain bulk copy, no arrival (the caller has posted the transaction bytes of the whole group)
__device__ __forceinline__ void bulk_g2s_elect(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n\t}" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_elect(uint64_t* bar, uint32_t bytes) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}" ::"r"(smem_u32(bar)), "r"(bytes)
      : "memory");
}

// kTrace: development build with event timestamps (a lane-0 region after every hand-off makes the
// issuing loops divergent for the compiler, so the production instantiation must not contain it)
template <bool kF16, bool kTrace, bool kDev>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
fused_pair_kernel(const __grid_constant__ PairParams P, const __grid_constant__ CUtensorMap x_map) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  PBars* bars = reinterpret_cast<PBars*>(smem + P.off_bar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + P.off_bar + sizeof(PBars));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  int trace_n = 0;
  (void)trace_n;
  uint32_t prof_acc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, prof_t = (uint32_t)clock();
  (void)prof_acc;
  (void)prof_t;
  const uint32_t rank = (uint32_t)blockIdx.x & 1u;      // == %cluster_ctarank for a cluster of 2 along x; 0 = leader
  const bool leader = rank == 0;
  if ((smem_u32(smem) & 1023u) != 0u || cluster_ctarank() != rank) {
    if (threadIdx.x == 0) printf("rafiki_b200: pair kernel: misaligned shared memory or unexpected cluster rank\n");
    __trap();
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < MAX_RAW_STAGES; ++s) {
      mbar_init(&bars->raw_full[s], 1);
      mbar_init(&bars->raw_empty[s], kConvWarps / 2);    // the four warps of the group that owns the stage
    }
    for (int s = 0; s < A_STAGES; ++s) {
      mbar_init(&bars->a_full[s], kConvWarps);       // 4 warps of the owning group x 2 CTAs
      mbar_init(&bars->a_free[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->acc1_full[b], 1);
      mbar_init(&bars->acc1_empty[b], 2 * kEpiWarps);
    }
    mbar_init(&bars->hid_ready, 2 * kEpiWarps);
    mbar_init(&bars->acc2_full, 1);
    mbar_init(&bars->acc3_full, 1);
    mbar_init(&bars->w_full, 1);
    mbar_init(&bars->w_free, kEpiWarps);
    mbar_init(&bars->w_ready, 2);
    fence_mbar_init();
  }
  if (warp == WARP_LOAD && lane == 0) tma_prefetch_desc(&x_map);
  if (warp == WARP_MMA) tmem_alloc_pair<TMEM_COLS>(tmem_slot);
  tc_fence_before_sync();
  cluster_sync_all();            // barriers of BOTH CTAs are initialised before anyone signals across the pair
  tc_fence_after_sync();
  if (*tmem_slot != 0u) {
    if (threadIdx.x == 0) printf("rafiki_b200: unexpected TMEM base %u\n", *tmem_slot);
    __trap();
  }
  constexpr uint32_t tmem_base = 0u;
  const uint32_t smem_base = smem_u32(smem);
  const int n_kb = P.n_kb;
  // Launched as a programmatic dependent of the previous kernel of the stream: everything up to here touched
  // shared and tensor memory only and may overlap that kernel's tail; the next kernel may start ITS prologue
  // now.  Every warp waits for the previous grid before it touches global memory -- except the loader, which
  // first starts the copies of this pair's first trial (parameters are immutable once a trial is loaded).
  griddep_launch_dependents();
  if (warp != WARP_LOAD && !P.hook.no_grid_wait) griddep_wait();
  const int dbg = kDev ? P.debug : 0;      // development switches (timing experiments; results are not valid)

  // ---- this pair's items: a contiguous range of (trial, pair tile) in trial-major order
  const int pair = (int)blockIdx.x >> 1;
  const int i_lo = pair * P.items_q + (pair < P.items_r ? pair : P.items_r);
  const int n_items = P.items_q + (pair < P.items_r ? 1 : 0);
  // leader's barriers as seen through the cluster window
  const uint32_t l_a_full0 = map_to_cta(smem_u32(&bars->a_full[0]), 0);
  const uint32_t l_acc1_empty0 = map_to_cta(smem_u32(&bars->acc1_empty[0]), 0);
  const uint32_t l_hid_ready = map_to_cta(smem_u32(&bars->hid_ready), 0);
  const uint32_t l_w_ready = map_to_cta(smem_u32(&bars->w_ready), 0);

  if (warp < kEpiWarps) {
    // =============================================================== epilogue warps (own 128 rows)
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsEpi));
    const int quarter = warp & 3;
    const int half = warp >> 2;                                // 0: hidden columns 0-63 + softmax; 1: columns 64-127
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t hid_row_addr = lane_addr + COL_HID;
    const float* bias = reinterpret_cast<const float*>(smem + P.off_bias);
    uint32_t v3[16];
    bool have_prev = false;
    bool may_store = P.hook.n_wait == 0;      // output in a peer's memory: wait until the slots have been consumed
    int prev_t = 0, prev_pt = 0;
    // The probabilities of a finished item leave in two steps on different warps.  The four half-0 warps turn
    // the logits into probabilities and stage them densely in shared memory (finish_math, between the two hidden
    // epilogues of the NEXT item, while the tensor pipe works on its layer 2); the four half-1 warps, which have
    // nothing to do at the end of an item, copy the staged tile to global memory with 16-byte stores
    // (finish_store) -- off the item-to-item chain of the softmax warps.  Two named barriers hand the staging
    // buffer back and forth (arrive on one side, sync on the other, 128 + 128 threads).
    bool stage_used = false;
    auto finish_math = [&]() {
      float z[CPAD];
      const float inv = softmax16(v3, bias + 2 * HMAX, P.n_classes, z);
      PROF(8);
      const int C = P.n_classes;
      const uint32_t dst = smem_base + (uint32_t)P.off_out + (uint32_t)(row * C) * 4u;
      // A lone trial of a larger ensemble (one trial per GPU) stores float(p / total) with ONE rounding of the
      // exact quotient, which is what the single-CTA kernel's float64 division gives for at most 32 trials.
      // A power-of-two total is an exact scaling; otherwise one correctly rounded float division.  (The
      // float64 division itself costs ~10 us per launch on this path and must stay off it.)
      float scale = inv;
      if (P.div_mode == 1) scale = inv * P.div_recip;      // exact: a power of two
#pragma unroll
      for (int j = 0; j < CPAD; ++j) z[j] = P.div_mode == 2 ? __fdiv_rn(z[j] * inv, P.div_f) : z[j] * scale;
      if (stage_used) named_bar_sync(BAR_STAGE_FREE, 256);
      PROF(9);
      if (C == 10) {                                        // the reference's task: five 8-byte stores, no predicates
#pragma unroll
        for (int j = 0; j < 10; j += 2)
          asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(dst + 4u * j), "f"(z[j]), "f"(z[j + 1]) : "memory");
      } else {
#pragma unroll
        for (int j = 0; j < CPAD; ++j) sts_f32_if(dst + 4u * j, z[j], j < C);
      }
      __threadfence_block();
      named_bar_arrive(BAR_STAGE_FULL, 256);
      stage_used = true;
      PROF(10);
    };
    auto finish_store = [&]() {
      if (!may_store) {
        if (lane < P.hook.n_wait) {                 // one flag word per lane: the polls overlap
          const long long t0 = clock64();
          unsigned int v;
          do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(P.hook.wait_flags + lane) : "memory");
            if (clock64() - t0 > 4000000000LL) {
              printf("rafiki_b200: pair kernel timed out waiting for its output slot (block %d)\n", blockIdx.x);
              __trap();
            }
          } while ((int)(v - P.hook.wait_value) < 0);
        }
        __syncwarp();
        may_store = true;
      }
      named_bar_sync(BAR_STAGE_FULL, 256);
      PROF(8);
      {
        const int C = P.n_classes;
        const float* stage = reinterpret_cast<const float*>(smem + P.off_out);
        const long long first = (long long)prev_pt * (2 * TILE_M) + (long long)rank * TILE_M;
        long long rows = P.n - first;
        rows = rows < 0 ? 0 : (rows < TILE_M ? rows : TILE_M);
        float* obase = P.out + (long long)prev_t * P.trial_stride;
        if (P.hook.n_owner > 0) {                  // scatter form: this tile's rows belong to one owner GPU
          int j = 0;
          while (j + 1 < P.hook.n_owner && prev_pt >= P.hook.owner_bound[j + 1]) ++j;
          obase = P.hook.out_owner[j];
        }
        store_tile_probs(obase + first * C, stage, (int)rows * C, row);
      }
      named_bar_arrive(BAR_STAGE_FREE, 256);          // the loads of the staged tile are complete (their values were stored)
      PROF(9);
    };
    auto finish_prev = [&]() {
      if (dbg & 32) return;
      if (half == 0) finish_math(); else finish_store();
    };
    auto fetch_prev_logits = [&](uint32_t parity) {
      // both halves wait: the hidden tile may not be overwritten before the layer-3 MMAs have read it
      mbar_wait(&bars->acc3_full, parity);
      tc_fence_after_sync();
      if (half == 0) {
        tmem_ld16(lane_addr + COL_ACC3, v3);
        tmem_ld_wait();
      }
    };
    int epoch = 0, cur_t = -1;
    uint32_t acc2_phase = 0;
    ItemIter it;
    it.init(i_lo, P.n_pt);
    for (int item = 0; item < n_items; ++item, it.next(P.n_pt)) {
      const PairTrial& T = P.trial[it.t];
      const int buf = item & 1;
      const uint32_t par2 = (uint32_t)((item >> 1) & 1);
      if (it.t != cur_t) {
        // a new trial: finish the last item of the old one (its softmax reads the old b3), hand the weight
        // buffers back to the loader, wait for the new trial's parameters
        if (have_prev) {
          fetch_prev_logits((uint32_t)((item - 1) & 1));
          tc_fence_before_sync();
          finish_prev();
          have_prev = false;
        }
        if (epoch > 0) warp_arrive(&bars->w_free);
        mbar_wait(&bars->w_full, (uint32_t)(epoch & 1));
        ++epoch;
        cur_t = it.t;
      }
      // ---- layer 1: acc1[buf] -> hidden
      PROF(0);
      mbar_wait(&bars->acc1_full[buf], par2);
      PROF(1);
      if (warp == 0) PTRACE(3, 1);
      if (have_prev) fetch_prev_logits((uint32_t)((item - 1) & 1));
      PROF(2);
      if (warp == 0) PTRACE(3, 5);
      tc_fence_after_sync();
      {
        const int split = T.h1_pad < 64 ? T.h1_pad : 64;
        epilogue_hidden<kF16>(lane_addr + COL_ACC1 + buf * ACC_STRIDE, half ? split : 0, half ? T.h1_pad : split, bias,
                              kF16 ? kSubnormalScale : 1.0f, hid_row_addr);
      }
      tc_fence_before_sync();
      warp_arrive_leader(l_hid_ready);
      if (T.n_hidden != 2) warp_arrive_leader(l_acc1_empty0 + 8u * buf);
      PROF(3);
      if (warp == 0) PTRACE(3, 2);
      if (have_prev && half == 0) finish_prev();       // probabilities of the previous item: math and staging
      PROF(4);
      if (warp == 0) PTRACE(3, 6);
      if (T.n_hidden == 2) {
        mbar_wait(&bars->acc2_full, acc2_phase);
        PROF(5);
        if (warp == 0) PTRACE(3, 3);
        acc2_phase ^= 1u;
        tc_fence_after_sync();
        const int split = T.h2_pad < 64 ? T.h2_pad : 64;
        epilogue_hidden<kF16>(lane_addr + COL_ACC1 + buf * ACC_STRIDE, half ? split : 0, half ? T.h2_pad : split, bias + HMAX,
                              1.0f, hid_row_addr);
        tc_fence_before_sync();
        warp_arrive_leader(l_hid_ready);
        warp_arrive_leader(l_acc1_empty0 + 8u * buf);
        PROF(6);
        if (warp == 0) PTRACE(3, 4);
      }
      if (have_prev && half == 1) finish_prev();       // ... and their way out, on the warps that are idle now
      PROF(7);
      have_prev = true;
      prev_t = it.t;
      prev_pt = it.pt;
    }
    if (have_prev) {
      fetch_prev_logits((uint32_t)((n_items - 1) & 1));
      tc_fence_before_sync();
      finish_prev();
    }
  } else if (warp < kEpiWarps + kConvWarps) {
    // =============================================================== converter warps (own 128 rows)
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsConv));
    // A converter's loop is a serial chain (barrier probe, shared-memory loads, byte permutes, tensor-memory
    // stores and their wait, arrivals): its latency, not its instruction count, sets the pace.  So one warp
    // converts its 32 rows of a WHOLE raw stage per turn (128 bytes per row = two 64-wide K atoms = two A
    // stages, 8 independent 16-byte loads in flight), and the two groups of four warps take alternate raw
    // stages: the chain is paid once per two atoms and two chains overlap.
    const int cw = warp - kEpiWarps;
    const int quarter = cw & 3;
    const int par = cw >> 2;                                // this warp takes raw stages par, par + 2, ... of the pair's stream
    const int row = quarter * 32 + lane;
    const uint32_t raw_row = smem_base + (uint32_t)P.off_raw + (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const uint32_t a_dst0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + COL_A;
    uint32_t lane0 = lane == 0 ? 1u : 0u;
    asm volatile("" : "+r"(lane0));                         // keep the flag in a register (no S2R in the loop)
    int item = 0, kb = par;
    while (kb >= n_kb) { kb -= n_kb; ++item; }
    int rstage = par;                                       // raw_stages >= 3 > par
    uint32_t rphase = 0;
    auto convert32 = [&](uint32_t src, uint32_t c0, uint32_t dst) {      // 32 bytes of the row -> 16 columns
      const uint4 r0 = lds_u4(src + ((c0 ^ sw) << 4)), r1 = lds_u4(src + (((c0 + 1u) ^ sw) << 4));
      uint32_t q[16];
      q[0] = u8x2_to_op<kF16>(r0.x, 0);  q[1] = u8x2_to_op<kF16>(r0.x, 2);
      q[2] = u8x2_to_op<kF16>(r0.y, 0);  q[3] = u8x2_to_op<kF16>(r0.y, 2);
      q[4] = u8x2_to_op<kF16>(r0.z, 0);  q[5] = u8x2_to_op<kF16>(r0.z, 2);
      q[6] = u8x2_to_op<kF16>(r0.w, 0);  q[7] = u8x2_to_op<kF16>(r0.w, 2);
      q[8] = u8x2_to_op<kF16>(r1.x, 0);  q[9] = u8x2_to_op<kF16>(r1.x, 2);
      q[10] = u8x2_to_op<kF16>(r1.y, 0); q[11] = u8x2_to_op<kF16>(r1.y, 2);
      q[12] = u8x2_to_op<kF16>(r1.z, 0); q[13] = u8x2_to_op<kF16>(r1.z, 2);
      q[14] = u8x2_to_op<kF16>(r1.w, 0); q[15] = u8x2_to_op<kF16>(r1.w, 2);
      tmem_st16(dst, q);
    };
    while (item < n_items) {
      const int atom0 = 2 * kb;                             // always < n_atoms
      const bool two = atom0 + 1 < P.n_atoms;               // the ragged end of the row may leave one atom only
      const int ga = item * P.n_atoms + atom0;              // global atom counter of this pair
      const int as0 = ga % A_STAGES, round0 = ga / A_STAGES;
      const int as1 = as0 + 1 == A_STAGES ? 0 : as0 + 1, round1 = as0 + 1 == A_STAGES ? round0 + 1 : round0;
      PROF(0);
      if (!(dbg & 8)) mbar_wait(&bars->raw_full[rstage], rphase);
      PROF(1);
      if (warp == kEpiWarps) PTRACE(0, 1);
      // The two atoms are handed over one at a time: the first goes to the issuer as soon as ITS stage was free
      // and its 64 bytes are in tensor memory, while the second may still be waiting for a stage to come back
      // (stages return one MMA group, ~256 cycles, apart).
      if (round0 > 0) mbar_wait(&bars->a_free[as0], (uint32_t)((round0 - 1) & 1));
      PROF(3);
      if (warp == kEpiWarps) PTRACE(0, 2);
      tc_fence_after_sync();
      const uint32_t rsrc = raw_row + (uint32_t)rstage * RAW_STAGE;
      const uint32_t dst0 = a_dst0 + (uint32_t)as0 * A_COLS, dst1 = a_dst0 + (uint32_t)as1 * A_COLS;
      const int left = P.d - atom0 * ATOM_K;                // K elements of this row from here on (> 0)
      const int live = left >= 2 * ATOM_K ? 8 : (left + 15) >> 4;     // 16-byte chunks the MMAs read: 1 .. 8
      convert32(rsrc, 0u, dst0);
      if (live > 2) convert32(rsrc, 2u, dst0 + 16u);
      tmem_st_wait();
      tc_fence_before_sync();
      __syncwarp();
      if (lane0) mbar_arrive_cluster(l_a_full0 + 8u * as0);
      PROF(2);
      if (two) {
        if (round1 > 0) mbar_wait(&bars->a_free[as1], (uint32_t)((round1 - 1) & 1));
        tc_fence_after_sync();
        if (live > 4) convert32(rsrc, 4u, dst1);
        if (live > 6) convert32(rsrc, 6u, dst1 + 16u);
      }
      __syncwarp();
      if (lane0 && !(dbg & 8)) mbar_arrive(&bars->raw_empty[rstage]);      // the bytes are in registers
      if (two) {
        tmem_st_wait();
        PROF(4);
        tc_fence_before_sync();
        __syncwarp();
        if (lane0) mbar_arrive_cluster(l_a_full0 + 8u * as1);
      }
      PROF(5);
      if (warp == kEpiWarps) PTRACE(0, 3);
      kb += 2;
      if (kb >= n_kb) { kb -= n_kb; ++item; if (kb >= n_kb) { kb -= n_kb; ++item; } }
      rstage += 2;
      if (rstage >= P.raw_stages) { rstage -= P.raw_stages; rphase ^= 1u; }
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsLight));
  }
  if (warp == WARP_LOAD) {
    // =============================================================== loader (own CTA's shared memory)
    // At a trial change: this CTA's half of W1 / W2 / W3 and the biases, one transaction group on
    // w_full, once the epilogue warps have released the buffers.  Per raw stage: the uint8 tile of
    // this CTA's 128 rows, 128 columns.
    int rstage = 0, epoch = 0, cur_t = -1;
    uint32_t rphase = 0;
    ItemIter it;
    it.init(i_lo, P.n_pt);
    for (int item = 0; item < n_items; ++item, it.next(P.n_pt)) {
      if (it.t != cur_t) {
        const PairTrial& T = P.trial[it.t];
        if (epoch > 0) mbar_wait(&bars->w_free, (uint32_t)((epoch - 1) & 1));
        const uint32_t w1_half = (uint32_t)(T.h1_pad >> 1) * 128u;               // bytes of one atom's half
        const int hl_pad = T.n_hidden == 2 ? T.h2_pad : T.h1_pad;
        const int atoms2 = T.n_hidden == 2 ? (T.h1_pad + 63) / 64 : 0, atoms3 = (hl_pad + 63) / 64;
        const uint32_t w2_half = (uint32_t)(T.h2_pad >> 1) * 128u;
        const uint32_t total = (uint32_t)P.n_atoms * w1_half + (uint32_t)atoms2 * w2_half + (uint32_t)atoms3 * W3_ATOM +
                               BIAS_BYTES;
        mbar_expect_tx_elect(&bars->w_full, total);
        for (int a = 0; a < P.n_atoms; ++a)
          bulk_g2s_elect(smem + a * W1_ATOM, T.w1 + (size_t)a * 2 * w1_half + (size_t)rank * w1_half, w1_half, &bars->w_full);
        for (int a = 0; a < atoms2; ++a)
          bulk_g2s_elect(smem + P.off_w2 + a * (HHALF * 128), T.w2 + (size_t)a * 2 * w2_half + (size_t)rank * w2_half,
                         w2_half, &bars->w_full);
        for (int a = 0; a < atoms3; ++a)
          bulk_g2s_elect(smem + P.off_w3 + a * W3_ATOM, T.w3 + (size_t)a * 2 * W3_ATOM + (size_t)rank * W3_ATOM, W3_ATOM,
                         &bars->w_full);
        bulk_g2s_elect(smem + P.off_bias, T.bias, BIAS_BYTES, &bars->w_full);
        if (epoch == 0 && !P.hook.no_grid_wait) griddep_wait();        // the queries (and everything after) may come from the previous kernel
        ++epoch;
        cur_t = it.t;
      }
      const int row0 = it.pt * (2 * TILE_M) + (int)rank * TILE_M;
      for (int kb = 0; kb < n_kb && !(dbg & 8); ++kb) {
        PROF(0);
        mbar_wait(&bars->raw_empty[rstage], rphase ^ 1u);
        PROF(1);
        tma_load_2d_expect_elect(smem + P.off_raw + rstage * RAW_STAGE, &x_map, kb * KB, row0, RAW_STAGE,
                                 &bars->raw_full[rstage]);
        if (++rstage == P.raw_stages) { rstage = 0; rphase ^= 1u; }
      }
    }
  } else if (warp == WARP_FWD) {
    // =============================================================== "weights landed" -> the leader
    int epoch = 0, cur_t = -1;
    ItemIter it;
    it.init(i_lo, P.n_pt);
    for (int item = 0; item < n_items; ++item, it.next(P.n_pt)) {
      if (it.t != cur_t) {
        mbar_wait(&bars->w_full, (uint32_t)(epoch & 1));
        warp_arrive_leader(l_w_ready);
        ++epoch;
        cur_t = it.t;
      }
    }
    if (kTrace && P.trace != nullptr && blockIdx.x == 0) {
      // development: this idle warp watches the A stages come back (exact arrival times of a_free)
      int as = 0;
      uint32_t ph = 0;
      for (int n = 0; n < n_items * P.n_atoms; ++n) {
        mbar_wait(&bars->a_free[as], ph);
        PTRACE(1, 2);
        if (++as == A_STAGES) { as = 0; ph ^= 1u; }
      }
    }
    // =============================================================== carried cross-GPU reduction (see PeerHook)
    // This warp is idle for the rest of the launch.  In the trial-per-GPU deployment it adds, for the rows
    // this GPU owns, the partial means every rank stored here during the PREVIOUS step (float64, rank order,
    // one rounding) and stores them into the root's result: the collective rides inside the next trial
    // kernel instead of standing between two of them.
    if (P.hook.n_carry > 0) {
      const int G = P.hook.n_carry;
      if (lane < G) {
        const long long t0 = clock64();
        unsigned int v;
        do {
          asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(P.hook.carry_arrived + lane) : "memory");
          if (clock64() - t0 > 4000000000LL) {
            printf("rafiki_b200: carried reduction timed out waiting for rank %d (block %d)\n", lane, blockIdx.x);
            __trap();
          }
        } while ((int)(v - P.hook.carry_step) < 0);
      }
      __syncwarp();
      const long long span = P.hook.carry_v_hi - P.hook.carry_v_lo;
      const long long per = (span + gridDim.x - 1) / gridDim.x;
      const long long lo = P.hook.carry_v_lo + per * blockIdx.x;
      const long long hi = lo + per < P.hook.carry_v_hi ? lo + per : P.hook.carry_v_hi;
      // Eight 16-byte loads in flight per lane (kPos positions x up to kSlots slots per round): the loop is bound
      // by memory latency (the slots were written by peers, the trial kernel keeps L2 busy), not by the adds
      auto carry_loop = [&](auto kPosTag, auto kSlotsTag) {
        constexpr int kPos = decltype(kPosTag)::value, kSlots = decltype(kSlotsTag)::value;
        for (long long i = lo + lane; i < hi; i += 32 * kPos) {
          double acc[kPos][4];
#pragma unroll
          for (int u = 0; u < kPos; ++u) acc[u][0] = acc[u][1] = acc[u][2] = acc[u][3] = 0.0;
          for (int j0 = 0; j0 < G; j0 += kSlots) {
            float4 v[kPos][kSlots];
#pragma unroll
            for (int q = 0; q < kSlots; ++q)
              if (j0 + q < G) {
                const float4* sl = reinterpret_cast<const float4*>(P.hook.carry_slot[j0 + q]);
#pragma unroll
                for (int u = 0; u < kPos; ++u)
                  if (i + 32 * u < hi) v[u][q] = __ldcg(sl + i + 32 * u);
              }
#pragma unroll
            for (int q = 0; q < kSlots; ++q)
              if (j0 + q < G) {
#pragma unroll
                for (int u = 0; u < kPos; ++u)
                  if (i + 32 * u < hi) {
                    acc[u][0] += (double)v[u][q].x; acc[u][1] += (double)v[u][q].y;
                    acc[u][2] += (double)v[u][q].z; acc[u][3] += (double)v[u][q].w;
                  }
              }
          }
#pragma unroll
          for (int u = 0; u < kPos; ++u)
            if (i + 32 * u < hi) {
              float4 o;
              o.x = (float)acc[u][0]; o.y = (float)acc[u][1]; o.z = (float)acc[u][2]; o.w = (float)acc[u][3];
              reinterpret_cast<float4*>(P.hook.carry_out)[i + 32 * u] = o;
            }
        }
      };
      if (G <= 2)
        carry_loop(std::integral_constant<int, 4>{}, std::integral_constant<int, 2>{});
      else
        carry_loop(std::integral_constant<int, 2>{}, std::integral_constant<int, 4>{});
      __syncwarp();      // the sums are published at the end of the kernel, with the output hand-shake
    }
  } else if (warp == WARP_MMA && leader) {
    // =============================================================== layer-1 MMA issuer (leader CTA, for the pair)
    // A: tensor memory of each CTA (its own 128 rows); B: each CTA's half of W1, resident; D: tensor
    // memory of each CTA.  One commit per atom hands the A stage back to the converters of both CTAs.
    // The loop is a serial chain on the critical path (per atom: barrier probe, four MMAs, a commit); it has
    // to turn in fewer cycles than the four MMAs take, so the common case (a whole 64-wide atom) is kept short.
    {
      const uint32_t w1_lo0 = desc_lo_sw128(smem_base);
      const int n_whole = P.d >> 6;                        // atoms with all four K steps
      int as = 0, epoch = 0, cur_t = -1;
      uint32_t aphase = 0;
      ItemIter it;
      it.init(i_lo, P.n_pt);
      for (int item = 0; item < n_items; ++item, it.next(P.n_pt)) {
        const PairTrial& T = P.trial[it.t];
        if (it.t != cur_t) {
          mbar_wait(&bars->w_ready, (uint32_t)(epoch & 1));
          ++epoch;
          cur_t = it.t;
        }
        const int buf = item & 1;
        const uint32_t idesc1 = idesc_16bit_f32(kF16, 2 * TILE_M, T.h1_pad);
        const uint32_t d_tmem = tmem_base + COL_ACC1 + buf * ACC_STRIDE;
        PROF(0);
        mbar_wait(&bars->acc1_empty[buf], (uint32_t)(((item >> 1) & 1) ^ 1));
        PROF(1);
        PTRACE(2, 4);
        uint32_t b = w1_lo0;
        for (int atom = 0; atom < P.n_atoms; ++atom, b += (uint32_t)(W1_ATOM >> 4)) {
          mbar_wait(&bars->a_full[as], aphase);
          PROF(2);
          PTRACE(2, 1);
          tc_fence_after_sync();
          const uint32_t a = tmem_base + COL_A + (uint32_t)as * A_COLS;
          if (!(dbg & 4)) {
            if (atom < n_whole)
              mma4_f16_ts_pair_whole(d_tmem, a, b, idesc1, atom != 0);
            else
              mma4_f16_ts_pair_elect(d_tmem, a, desc_from_lo(b), idesc1, atom != 0, (P.d & (ATOM_K - 1)) >> 4);
          }
          PROF(3);
          PTRACE(2, 2);
          mma_commit_pair_elect(&bars->a_free[as]);
          PROF(4);
          if (++as == A_STAGES) { as = 0; aphase ^= 1u; }
        }
        mma_commit_pair_elect(&bars->acc1_full[buf]);
      }
    }
  } else if (warp == WARP_TAIL && leader) {
    // =============================================================== layers 2 and 3 (leader CTA, for the pair)
    const uint32_t w2_lo = desc_lo_sw128(smem_base + (uint32_t)P.off_w2), w3_lo = desc_lo_sw128(smem_base + (uint32_t)P.off_w3);
    auto issue_small_layer = [&](uint32_t d_tmem, uint32_t b_lo, uint32_t b_atom, uint32_t idesc, int ksteps) {
      for (int k0 = 0; k0 < ksteps; k0 += 4) {
        const uint32_t a = tmem_base + COL_HID + (uint32_t)k0 * 8u;
        const uint32_t b = b_lo + (uint32_t)(k0 >> 2) * b_atom;
        mma4_f16_ts_pair_elect(d_tmem, a, desc_from_lo(b), idesc, k0 != 0, ksteps - k0 < 4 ? ksteps - k0 : 4);
      }
    };
    int epoch = 0, cur_t = -1;
    uint32_t hid_phase = 0;
    ItemIter it;
    it.init(i_lo, P.n_pt);
    for (int item = 0; item < n_items; ++item, it.next(P.n_pt)) {
      const PairTrial& T = P.trial[it.t];
      if (it.t != cur_t) {
        mbar_wait(&bars->w_ready, (uint32_t)(epoch & 1));
        ++epoch;
        cur_t = it.t;
      }
      const int buf = item & 1;
      if (T.n_hidden == 2) {
        PROF(0);
        mbar_wait(&bars->hid_ready, hid_phase);
        PROF(1);
        PTRACE(4, 1);
        hid_phase ^= 1u;
        tc_fence_after_sync();
        issue_small_layer(tmem_base + COL_ACC1 + (uint32_t)buf * ACC_STRIDE, w2_lo, (HHALF * 128u) >> 4,
                          idesc_16bit_f32(kF16, 2 * TILE_M, T.h2_pad), T.h1_pad >> 4);
        mma_commit_pair_elect(&bars->acc2_full);
        PROF(2);
        PTRACE(4, 2);
      }
      mbar_wait(&bars->hid_ready, hid_phase);
      PROF(3);
      PTRACE(4, 3);
      hid_phase ^= 1u;
      tc_fence_after_sync();
      const int hl_pad = T.n_hidden == 2 ? T.h2_pad : T.h1_pad;
      issue_small_layer(tmem_base + COL_ACC3, w3_lo, (uint32_t)W3_ATOM >> 4, idesc_16bit_f32(kF16, 2 * TILE_M, CPAD), hl_pad >> 4);
      mma_commit_pair_elect(&bars->acc3_full);
      PROF(4);
      PTRACE(4, 4);
    }
  }
  if (kDev && P.prof != nullptr && blockIdx.x < 2 && lane == 0) {
    PROF(7);
    for (int k = 0; k < 12; ++k) P.prof[((int)blockIdx.x * 20 + warp) * 12 + k] = prof_acc[k];
  }
  __syncwarp();
  tc_fence_before_sync();
  cluster_sync_all();            // neither CTA leaves (or frees tensor memory) while the other may still signal it
  if (warp == WARP_MMA) tmem_dealloc_pair<TMEM_COLS>(tmem_base);
  if ((P.hook.n_signal > 0 || P.hook.n_carry > 0) && threadIdx.x == 0) {
    // Every store of this CTA (its tiles' outputs, its share of the carried sums) precedes the cluster barrier
    // above.  ONE system-scope fence per CTA, then the last CTA out publishes the flags with plain stores: a
    // release store per flag would pay a system-wide memory barrier (~1.5 us) each.
    __threadfence_system();
    if (atomicAdd(P.hook.counter, 1u) == gridDim.x - 1) {
      atomicExch(P.hook.counter, 0u);
      __threadfence();
      for (int q = 0; q < P.hook.n_signal; ++q)
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(P.hook.signal_flags[q]), "r"(P.hook.signal_value) : "memory");
      if (P.hook.n_carry > 0) {
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(P.hook.carry_done), "r"(P.hook.carry_step) : "memory");
        for (int j = 0; j < P.hook.n_carry; ++j)
          asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(P.hook.carry_consumed[j]), "r"(P.hook.carry_step)
                       : "memory");
      }
    }
  }
}

}  // namespace

// ------------------------------------------------------------------------------- host side
namespace rfk {

namespace {
thread_local PeerHook g_hook;
thread_local bool g_hook_set = false;
}  // namespace
void set_peer_hook(const PeerHook* hook) {
  g_hook_set = hook != nullptr;
  if (hook != nullptr) g_hook = *hook;
}
bool peer_hook_pending() { return g_hook_set; }

namespace {
constexpr int kMinPairTiles = 37;     // below this the single-CTA kernel's finer tiles fill the GPU better

struct PairLayout {
  int off_w2, off_w3, off_raw, raw_stages, off_bias, off_out, off_bar, total;
};
bool pair_layout(int d, PairLayout* L) {
  const int n_atoms = (d + ATOM_K - 1) / ATOM_K;
  L->off_w2 = n_atoms * W1_ATOM;
  L->off_w3 = L->off_w2 + W2_BYTES;
  L->off_raw = L->off_w3 + W3_BYTES;
  int stages = (SMEM_MAX - TAIL_BYTES - L->off_raw) / RAW_STAGE;
  if (stages < MIN_RAW_STAGES) return false;
  if (stages > MAX_RAW_STAGES) stages = MAX_RAW_STAGES;
  L->raw_stages = stages;
  L->off_bias = L->off_raw + stages * RAW_STAGE;
  L->off_out = L->off_bias + (BIAS_BYTES + 15) / 16 * 16;
  L->off_bar = L->off_out + OUT_BYTES;
  L->total = L->off_bias + TAIL_BYTES;
  return true;
}
}  // namespace

// The pair kernel takes a batch when it is large enough to give every pair work and the trial's
// half-weights plus a raw ring fit in shared memory.  The choice depends on n and d only, never on
// the number of trials, so a trial alone and the same trial inside an ensemble run the same kernel.
bool pair_applicable(int d, int64_t n) {
  PairLayout L;
  if (getenv("RFK_NO_PAIR") != nullptr) return false;
  const int64_t n_pt = (n + 2 * TILE_M - 1) / (2 * TILE_M);
  return n_pt >= kMinPairTiles && n_pt < (1ll << 22) && pair_layout(d, &L);
}

size_t pair_workspace_bytes(int w, int n_classes, int64_t n) {
  return w > 1 ? (size_t)w * (size_t)n * (size_t)n_classes * sizeof(float) : 0;
}

int pair_forward(rfk_mlp* const* trials, int w, int total_trials, const uint8_t* x, int64_t n, float* out, void* ws,
                 size_t ws_bytes, cudaStream_t st) {
  RFK_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "fused forward: x must be 16-byte aligned");
  const int d = trials[0]->d;
  PairLayout L;
  RFK_REQUIRE(pair_layout(d, &L), "pair kernel: d = %d does not fit in shared memory", d);
  RFK_REQUIRE(w == 1 || (ws != nullptr && ws_bytes >= pair_workspace_bytes(w, trials[0]->n_classes, n)),
              "pair kernel: workspace of %zu bytes needed", pair_workspace_bytes(w, trials[0]->n_classes, n));
  CUtensorMap x_map;
  {
    const int rc = make_tmap_u8_2d(&x_map, x, (uint64_t)n, (uint64_t)d, (uint64_t)d, TILE_M, KB, true);
    if (rc != RFK_OK) return rc;
  }
  PairParams P;
  memset(&P, 0, sizeof(P));
  for (int t = 0; t < w; ++t) {
    FusedImages im;
    RFK_REQUIRE(fused_images(trials[t], &im), "pair kernel: trial %d was not packed for the fused path", t);
    P.trial[t].w1 = im.w1;
    P.trial[t].w2 = im.w2;
    P.trial[t].w3 = im.w3;
    P.trial[t].bias = im.bias;
    P.trial[t].h1_pad = im.h1_pad;
    P.trial[t].h2_pad = im.h2_pad;
    P.trial[t].n_hidden = im.n_hidden;
  }
  const int C = trials[0]->n_classes;
  P.n = n;
  P.d = d;
  P.n_kb = (d + KB - 1) / KB;
  P.n_atoms = (d + ATOM_K - 1) / ATOM_K;
  P.w = w;
  P.n_classes = C;
  P.n_pt = (int)((n + 2 * TILE_M - 1) / (2 * TILE_M));
  P.raw_stages = L.raw_stages;
  P.off_w2 = L.off_w2;
  P.off_w3 = L.off_w3;
  P.off_raw = L.off_raw;
  P.off_bias = L.off_bias;
  P.off_out = L.off_out;
  P.off_bar = L.off_bar;
  if (w == 1) {
    P.out = out;
    P.trial_stride = 0;
    P.divisor = (double)total_trials;
    P.div_f = (float)total_trials;
    P.div_mode = total_trials == 1 ? 0 : ((total_trials & (total_trials - 1)) == 0 ? 1 : 2);
    P.div_recip = 1.0f / (float)total_trials;
    if (g_hook_set) {          // the kernel itself waits for / publishes the peer hand-shake
      P.hook = g_hook;
      g_hook_set = false;
    }
  } else {
    P.out = static_cast<float*>(ws);
    P.trial_stride = (long long)n * C;
    P.divisor = 1.0;
    P.div_mode = 0;
    if (g_hook_set) {          // several trials per GPU in the peer step: carried sum only, the mean is the caller's
      P.hook = g_hook;
      g_hook_set = false;
      RFK_REQUIRE(P.hook.n_owner == 0 && P.hook.n_wait == 0 && P.hook.n_signal == 0 && P.hook.skip_mean == 1,
                  "pair kernel: a hook for several trials may only carry the previous step's sum");
    }
  }
  {
    const char* dbg = getenv("RFK_PAIR_DEBUG");
    P.debug = dbg ? atoi(dbg) : 0;
  }
  const bool f16 = trials[0]->dtype == RFK_F16;
  const char* prof_path = getenv("RFK_PAIR_PROF");
  unsigned int* dprof = nullptr;
  if (prof_path != nullptr && f16) {       // development aid: allocates and synchronises
    RFK_CUDA_TRY(cudaMalloc(&dprof, 2 * 20 * 12 * sizeof(unsigned int)));
    RFK_CUDA_TRY(cudaMemsetAsync(dprof, 0, 2 * 20 * 12 * sizeof(unsigned int), st));
    P.prof = dprof;
  }
  const char* trace_path = getenv("RFK_PAIR_TRACE");
  long long* dtrace = nullptr;
  const size_t trace_elems = 2 * 5 * kTraceLen;
  if (trace_path != nullptr && f16) {      // development aid: allocates and synchronises
    RFK_CUDA_TRY(cudaMalloc(&dtrace, trace_elems * sizeof(long long)));
    RFK_CUDA_TRY(cudaMemsetAsync(dtrace, 0, trace_elems * sizeof(long long), st));
    P.trace = dtrace;
  }
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  RFK_REQUIRE(dev >= 0 && dev < 64, "pair kernel: device ordinal %d not supported", dev);
  static std::atomic<bool> attr_set[64];
  if (!attr_set[dev].load(std::memory_order_acquire)) {     // per device; setting it twice is harmless
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_pair_kernel<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_pair_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_pair_kernel<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
    cudaFuncAttributes fa;
    RFK_CUDA_TRY(cudaFuncSetAttribute(fused_pair_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
    RFK_CUDA_TRY(cudaFuncGetAttributes(&fa, fused_pair_kernel<true, false, false>));
    if (fa.numRegs * (kEpiWarps + kConvWarps + 4) < kEpiWarps * kRegsEpi + kConvWarps * kRegsConv + 4 * kRegsLight) {
      set_error("pair kernel was compiled with %d registers per thread; its setmaxnreg budget needs %d", fa.numRegs,
                kRegsLaunch);
      return RFK_ERR_UNSUPPORTED;
    }
    attr_set[dev].store(true, std::memory_order_release);
  }
  const long long items = (long long)w * P.n_pt;
  int pairs = num_sms() / 2;
  if (items < pairs) pairs = (int)items;
  P.items_q = (int)(items / pairs);
  P.items_r = (int)(items % pairs);
  {
    // a programmatic dependent of whatever precedes it in the stream: its prologue and first weight copies
    // overlap the tail of that kernel (see griddep_wait in the kernel); RFK_NO_PDL=1 launches it ordinarily
    static const bool pdl = getenv("RFK_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(2 * pairs));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = (size_t)L.total;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t le;
    if (f16 && dtrace != nullptr)
      le = cudaLaunchKernelEx(&cfg, fused_pair_kernel<true, true, false>, P, x_map);
    else if (f16 && (P.debug != 0 || dprof != nullptr))
      le = cudaLaunchKernelEx(&cfg, fused_pair_kernel<true, false, true>, P, x_map);
    else if (f16)
      le = cudaLaunchKernelEx(&cfg, fused_pair_kernel<true, false, false>, P, x_map);
    else
      le = cudaLaunchKernelEx(&cfg, fused_pair_kernel<false, false, false>, P, x_map);
    RFK_CUDA_TRY(le);
  }
  RFK_LAUNCH_CHECK("fused_pair_kernel");
  if (dprof != nullptr) {
    unsigned int hp[2 * 20 * 12];
    RFK_CUDA_TRY(cudaMemcpyAsync(hp, dprof, sizeof(hp), cudaMemcpyDeviceToHost, st));
    RFK_CUDA_TRY(cudaStreamSynchronize(st));
    cudaFree(dprof);
    FILE* f = fopen(prof_path, "w");
    if (f) {
      for (int c = 0; c < 2; ++c)
        for (int wv = 0; wv < 20; ++wv) {
          fprintf(f, "%d %d", c, wv);
          for (int k = 0; k < 12; ++k) fprintf(f, " %u", hp[(c * 20 + wv) * 12 + k]);
          fprintf(f, "\n");
        }
      fclose(f);
    }
  }
  if (dtrace != nullptr) {
    std::vector<long long> hall(trace_elems);
    RFK_CUDA_TRY(cudaMemcpyAsync(hall.data(), dtrace, trace_elems * sizeof(long long), cudaMemcpyDeviceToHost, st));
    RFK_CUDA_TRY(cudaStreamSynchronize(st));
    cudaFree(dtrace);
    FILE* f = fopen(trace_path, "w");
    if (f) {
      static const char* roles[5] = {"conv", "load", "mma", "epi", "tail"};
      for (int c = 0; c < 2; ++c)
        for (int r = 0; r < 5; ++r)
          for (int i = 0; i < kTraceLen; ++i) {
            const long long v = hall[(c * 5 + r) * kTraceLen + i];
            if (v != 0) fprintf(f, "%d %s %d %lld %lld\n", c, roles[r], i, v >> 4, v & 15);
          }
      fclose(f);
    }
  }
  if (w > 1 && !P.hook.skip_mean) {
    const float* ptrs[RFK_MAX_TRIALS];
    for (int t = 0; t < w; ++t) ptrs[t] = static_cast<const float*>(ws) + (size_t)t * (size_t)n * C;
    // float64 sum in trial order, divided by the size of the WHOLE ensemble, one rounding
    return ensemble_partial_mean(ptrs, w, n, C, out, (double)total_trials, st);
  }
  return RFK_OK;
}

}  // namespace rfk
