// Kernel 2, layered form: decode -> [GEMM + bias + ReLU] x L -> GEMM + bias + softmax, for trials
// too wide for the fused kernel (BASELINE config 4: D 3072, three hidden layers of 2048 units).
//
// Replaces the Keras graph of TfFeedForward._build_model run by Model.predict
// (examples/models/image_classification/TfFeedForward.py:95,141-157) layer by layer.  Here the
// GEMMs are genuinely tensor-bound (29.4 Mflop per query), so each layer is a classic persistent
// tcgen05 GEMM:
//   C[M, N] = act(A[M, K] * B[N, K]^T + bias),   A, B 16-bit (fp16 / bf16), fp32 accumulation
//   * A (activations, row-major in HBM) arrives by TMA tensor loads with the 128-byte swizzle,
//     B (weights) is pre-packed at rfk_mlp_create into the exact shared-memory image of each
//     (N tile, K block) and arrives by one cp.async.bulk per stage; 4-stage mbarrier ring
//   * tcgen05.mma M 128, N 256 (N 16 for the softmax layer), K 16; two accumulator buffers in
//     TMEM so the epilogue of tile i overlaps the MMAs of tile i+1
//   * epilogue: TMEM -> registers, + bias, ReLU and 16-bit pack in one cvt, 16-byte global stores
//     (hidden layers); bias + softmax -> fp32 probabilities (last layer)
// Layer 0 reads the uint8 queries through kernel 1 (decode to 16-bit, BatchNormalization folded
// into the first layer's weights, see mlp.cu).
#include "pack.cuh"
#include "rfk_internal.cuh"
#include "sm100.cuh"
#include "tmap.cuh"

namespace {

using namespace sm100;

constexpr int BM = 128, BK = 64, STAGES = 4;
constexpr int A_STAGE = BM * 128;                    // 16 KB
constexpr int kEpiWarps = 8;
constexpr int kThreads = (4 + kEpiWarps) * 32;       // warps 0 loader, 1 MMA, 2-3 idle, 4-11 epilogue

template <int BN>
struct Cfg {
  static constexpr int B_STAGE = BN * 128;
  static constexpr int OFF_B = STAGES * A_STAGE;
  static constexpr int OFF_BAR = OFF_B + STAGES * B_STAGE;
  static constexpr int NUM_BARS = 2 * STAGES + 4;
  static constexpr int SMEM = OFF_BAR + NUM_BARS * 8 + 16 + 1024;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
};

struct GemmParams {
  const uint8_t* b_img;     // [n_tiles_n][n_kb][BN rows x 128 B] swizzled
  const float* bias;        // [n_tiles_n * BN] (zero padded)
  void* c;                  // hidden layers: 16-bit [M, ldc]; last layer: float [M, n_classes]
  long long m;
  int n_kb, n_tiles_n, n_tiles_m, ldc, n_classes;
};

struct Bars {
  uint64_t full[STAGES], empty[STAGES], acc_full[2], acc_empty[2];
};

template <bool kF16>
__device__ __forceinline__ uint32_t relu_pack(float lo, float hi) {
  uint32_t r;
  if (kF16)
    asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  else
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

template <int BN, bool kF16, bool kLast>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bias_act_kernel(const __grid_constant__ GemmParams P, const __grid_constant__ CUtensorMap a_map) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  Bars* bars = reinterpret_cast<Bars*>(smem + C::OFF_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + C::OFF_BAR + C::NUM_BARS * 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->acc_full[b], 1);
      mbar_init(&bars->acc_empty[b], kEpiWarps * 32);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) tma_prefetch_desc(&a_map);
  if (warp == 1) tmem_alloc<C::TMEM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t smem_base = smem_u32(smem);
  const int n_tiles = P.n_tiles_m * P.n_tiles_n;
  const int n_kb = P.n_kb;

  if (warp == 0) {
    // ------------------------------------------------------------ loader: A by TMA, B by bulk copy
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int tm = tile / P.n_tiles_n, tn = tile % P.n_tiles_n;
      const uint8_t* bsrc = P.b_img + (size_t)tn * n_kb * C::B_STAGE;
      for (int kb = 0; kb < n_kb; ++kb) {
        mbar_wait(&bars->empty[stage], phase ^ 1u);
        if (lane == 0) {
          mbar_arrive_expect_tx(&bars->full[stage], A_STAGE + C::B_STAGE);
          tma_load_2d(smem + stage * A_STAGE, &a_map, kb * BK, tm * BM, &bars->full[stage]);
          bulk_g2s(smem + C::OFF_B + stage * C::B_STAGE, bsrc + (size_t)kb * C::B_STAGE, C::B_STAGE, &bars->full[stage]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (convergent warp, elected lane)
    int stage = 0, it = 0;
    uint32_t phase = 0;
    const uint32_t a_lo0 = desc_lo_sw128(smem_base), b_lo0 = desc_lo_sw128(smem_base + C::OFF_B);
    const uint32_t idesc = idesc_16bit_f32(kF16, BM, BN);
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      mbar_wait(&bars->acc_empty[buf], (uint32_t)(((it >> 1) & 1) ^ 1));
      tc_fence_after_sync();
      const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
      for (int kb = 0; kb < n_kb; ++kb) {
        mbar_wait(&bars->full[stage], phase);
        tc_fence_after_sync();
        const uint32_t a = a_lo0 + (uint32_t)stage * (A_STAGE >> 4);
        const uint32_t b = b_lo0 + (uint32_t)stage * (C::B_STAGE >> 4);
        mma4_f16_ss_elect(d_tmem, desc_from_lo(a), desc_from_lo(b), idesc, kb != 0, 4);
        mma_commit_elect(&bars->empty[stage]);
        if (kb == n_kb - 1) mma_commit_elect(&bars->acc_full[buf]);
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue
    const int e = warp - 4;
    const int quarter = warp & 3;                  // TMEM lanes 32*quarter.. (hardware rule: warp % 4)
    const int half = e >> 2;                       // column half of the tile
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int tm = tile / P.n_tiles_n, tn = tile % P.n_tiles_n;
      const int buf = it & 1;
      const long long grow = (long long)tm * BM + row;
      mbar_wait(&bars->acc_full[buf], (uint32_t)((it >> 1) & 1));
      tc_fence_after_sync();
      if (kLast) {
        if (half == 0) {
          uint32_t v[16];
          tmem_ld16(lane_addr + (uint32_t)(buf * BN), v);
          tmem_ld_wait();
          float z[16];
          float mx = -INFINITY;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            z[j] = __uint_as_float(v[j]) + __ldg(P.bias + j);
            if (j < P.n_classes) mx = fmaxf(mx, z[j]);
          }
          float sum = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            z[j] = j < P.n_classes ? __expf(z[j] - mx) : 0.f;
            sum += z[j];
          }
          const float inv = 1.0f / sum;
          if (grow < P.m) {
            float* o = static_cast<float*>(P.c) + grow * P.n_classes;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < P.n_classes) o[j] = z[j] * inv;
          }
        }
      } else {
        constexpr int kColsPerHalf = BN / 2;
        const int c_lo = half * kColsPerHalf;
        uint16_t* crow = static_cast<uint16_t*>(P.c) + grow * P.ldc + (long long)tn * BN;
        const float* brow = P.bias + (long long)tn * BN;
#pragma unroll 1
        for (int c = c_lo; c < c_lo + kColsPerHalf; c += 16) {
          uint32_t v[16];
          tmem_ld16(lane_addr + (uint32_t)(buf * BN + c), v);
          tmem_ld_wait();
          uint32_t p[8];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 bb = __ldg(reinterpret_cast<const float4*>(brow + c) + j);
            p[2 * j] = relu_pack<kF16>(__uint_as_float(v[4 * j + 0]) + bb.x, __uint_as_float(v[4 * j + 1]) + bb.y);
            p[2 * j + 1] = relu_pack<kF16>(__uint_as_float(v[4 * j + 2]) + bb.z, __uint_as_float(v[4 * j + 3]) + bb.w);
          }
          if (grow < P.m && (long long)tn * BN + c < P.ldc) {
            uint4* dst = reinterpret_cast<uint4*>(crow + c);
            dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
            dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
          }
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&bars->acc_empty[buf]);
    }
  }
  __syncwarp();
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------- host side
constexpr int BN_HIDDEN = 256, BN_LAST = 16;

struct TcLayer {
  uint8_t* b_img = nullptr;
  float* bias = nullptr;
  int n_kb = 0, n_tiles_n = 0, bn = 0, ld_out = 0, k_pad = 0;
};
struct TcState {
  std::vector<TcLayer> layers;
};

template <typename T>
int upload_vec(T** dst, const std::vector<T>& v) {
  RFK_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(dst), v.size() * sizeof(T)));
  RFK_CUDA_TRY(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return RFK_OK;
}

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

template <int BN, bool kF16, bool kLast>
int launch_gemm(const GemmParams& P, const CUtensorMap& a_map, cudaStream_t st) {
  auto kern = gemm_bias_act_kernel<BN, kF16, kLast>;
  static bool attr[64] = {false};
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  if (!attr[dev & 63]) {
    RFK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM));
    attr[dev & 63] = true;
  }
  const int tiles = P.n_tiles_m * P.n_tiles_n;
  const int sms = rfk::num_sms();
  kern<<<tiles < sms ? tiles : sms, kThreads, Cfg<BN>::SMEM, st>>>(P, a_map);
  RFK_LAUNCH_CHECK("gemm_bias_act_kernel");
  return RFK_OK;
}

}  // namespace

namespace rfk {

int tc_prepare(rfk_mlp* h) {
  RFK_REQUIRE(h->n_classes <= BN_LAST, "tensor-core path: at most %d classes (got %d)", BN_LAST, h->n_classes);
  TcState* s = new TcState();
  h->tc = s;
  const int n_layers = (int)h->layers.size();
  s->layers.resize(n_layers);
  for (int i = 0; i < n_layers; ++i) {
    const Layer& l = h->layers[i];
    TcLayer& t = s->layers[i];
    const bool last = i == n_layers - 1;
    t.bn = last ? BN_LAST : BN_HIDDEN;
    t.k_pad = l.in_pad;                         // multiple of 64
    t.n_kb = l.in_pad / BK;
    t.n_tiles_n = (l.out + t.bn - 1) / t.bn;
    t.ld_out = (l.out + 63) / 64 * 64;          // = next layer's in_pad
    // per N tile: rows [tn*bn, tn*bn + bn) of the [out_pad, in_pad] weight matrix, zero padded
    std::vector<uint16_t> tile_rows((size_t)t.bn * l.in_pad);
    std::vector<uint8_t> img, all;
    std::vector<float> bias((size_t)t.n_tiles_n * t.bn, 0.f);
    for (int tn = 0; tn < t.n_tiles_n; ++tn) {
      std::fill(tile_rows.begin(), tile_rows.end(), 0);
      for (int r = 0; r < t.bn; ++r) {
        const int j = tn * t.bn + r;
        if (j < l.out_pad) {
          memcpy(&tile_rows[(size_t)r * l.in_pad], &l.w16_host[(size_t)j * l.in_pad], (size_t)l.in_pad * 2);
          bias[(size_t)tn * t.bn + r] = l.b_fold_host[j];
        }
      }
      pack_kmajor_sw128(tile_rows.data(), t.bn, l.in_pad, &img);
      all.insert(all.end(), img.begin(), img.end());
    }
    int rc = upload_vec(&t.b_img, all);
    if (rc == RFK_OK) rc = upload_vec(&t.bias, bias);
    if (rc != RFK_OK) return rc;
  }
  return RFK_OK;
}

void tc_release(rfk_mlp* h) {
  TcState* s = static_cast<TcState*>(h->tc);
  if (s == nullptr) return;
  for (TcLayer& t : s->layers) {
    cudaFree(t.b_img);
    cudaFree(t.bias);
  }
  delete s;
  h->tc = nullptr;
}

// workspace: decoded queries [n, k_pad0] 16-bit + two ping-pong activation buffers [n, widest ld_out] 16-bit
size_t tc_workspace_bytes(const rfk_mlp* h, int64_t n) {
  const TcState* s = static_cast<const TcState*>(h->tc);
  if (s == nullptr) return 0;
  int widest = 64;
  for (size_t i = 0; i + 1 < s->layers.size(); ++i) widest = s->layers[i].ld_out > widest ? s->layers[i].ld_out : widest;
  return align256((size_t)n * s->layers[0].k_pad * 2) + 2 * align256((size_t)n * widest * 2);
}

int tc_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes, cudaStream_t st) {
  const TcState* s = static_cast<const TcState*>(h->tc);
  RFK_REQUIRE(s != nullptr, "tensor-core forward: trial was not packed for the layered path");
  const size_t need = tc_workspace_bytes(h, n);
  RFK_REQUIRE(ws != nullptr && ws_bytes >= need, "mlp_forward(tensor core): workspace too small (%zu < %zu)", ws_bytes,
              need);
  RFK_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "mlp_forward(tensor core): workspace must be 256-byte aligned");
  const bool f16 = h->dtype == RFK_F16;
  int widest = 64;
  for (size_t i = 0; i + 1 < s->layers.size(); ++i) widest = s->layers[i].ld_out > widest ? s->layers[i].ld_out : widest;
  char* base = static_cast<char*>(ws);
  char* dec = base;
  char* act[2] = {base + align256((size_t)n * s->layers[0].k_pad * 2),
                  base + align256((size_t)n * s->layers[0].k_pad * 2) + align256((size_t)n * widest * 2)};
  // kernel 1: uint8 -> 16-bit, exact; BatchNormalization lives in the first layer's weights
  int rc = rfk_decode_u8(x, dec, n, h->d, s->layers[0].k_pad, nullptr, nullptr, f16 ? RFK_F16 : RFK_BF16, st);
  if (rc != RFK_OK) return rc;
  const void* a = dec;
  int lda = s->layers[0].k_pad;
  const int n_layers = (int)s->layers.size();
  for (int i = 0; i < n_layers; ++i) {
    const TcLayer& t = s->layers[i];
    const bool last = i == n_layers - 1;
    CUtensorMap a_map;
    rc = make_tmap_16bit_sw128(&a_map, a, (uint64_t)n, (uint64_t)lda, (uint64_t)lda * 2, BM);
    if (rc != RFK_OK) return rc;
    GemmParams P;
    P.b_img = t.b_img;
    P.bias = t.bias;
    P.m = n;
    P.n_kb = t.n_kb;
    P.n_tiles_n = t.n_tiles_n;
    P.n_tiles_m = (int)((n + BM - 1) / BM);
    P.n_classes = h->n_classes;
    if (last) {
      P.c = probs;
      P.ldc = h->n_classes;
      rc = f16 ? launch_gemm<BN_LAST, true, true>(P, a_map, st) : launch_gemm<BN_LAST, false, true>(P, a_map, st);
    } else {
      P.c = act[i & 1];
      P.ldc = t.ld_out;
      rc = f16 ? launch_gemm<BN_HIDDEN, true, false>(P, a_map, st) : launch_gemm<BN_HIDDEN, false, false>(P, a_map, st);
      a = act[i & 1];
      lda = t.ld_out;
    }
    if (rc != RFK_OK) return rc;
  }
  return RFK_OK;
}

}  // namespace rfk
