// fp32 parity path of kernel 2: the trial network on CUDA cores, true fp32 FMA.
//
// Replaces the Keras graph of TfFeedForward (TfFeedForward.py:141-157) when the caller
// asks for RFK_F32: results stay within 1e-5 of the fp32 oracle (the tensor-core paths
// round operands to bf16).  Layer 0 reads the uint8 queries directly and applies the
// BatchNormalization affine while staging its A tile, so no decoded copy of the batch is
// written.  The softmax is one warp-slice per row with shuffle reductions.
#include "rfk_common.cuh"

namespace {

constexpr int BM = 128, BN = 64, BK = 16, TM = 8, TN = 4;
constexpr int kThreads = (BM / TM) * (BN / TN);  // 256

// C[M,N] = act(A'[M,K] @ W[K,N] + bias), A' = A*scale+shift when A is uint8.
template <bool kAU8>
__global__ void __launch_bounds__(kThreads)
sgemm_bias_act_kernel(const void* __restrict__ A_, const float* __restrict__ W,
                      const float* __restrict__ bias, float* __restrict__ C, int M, int N, int K,
                      const float* __restrict__ scale, const float* __restrict__ shift, int relu) {
  __shared__ float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  // A staging: thread -> (row = tid / 2, 8 consecutive k); B staging: (k = tid / 16, 4 consecutive n)
  const int a_row = tid >> 1, a_k = (tid & 1) * 8;
  const int b_k = tid >> 4, b_n = (tid & 15) * 4;

  for (int k0 = 0; k0 < K; k0 += BK) {
    {
      const int gm = m0 + a_row;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int gk = k0 + a_k + i;
        float v = 0.f;
        if (gm < M && gk < K) {
          if (kAU8) {
            v = (float)static_cast<const uint8_t*>(A_)[(int64_t)gm * K + gk];
            v = fmaf(v, __ldg(scale + gk), __ldg(shift + gk));
          } else {
            v = static_cast<const float*>(A_)[(int64_t)gm * K + gk];
          }
        }
        As[a_k + i][a_row] = v;
      }
      const int gk = k0 + b_k;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int gn = n0 + b_n + j;
        Bs[b_k][b_n + j] = (gk < K && gn < N) ? __ldg(W + (int64_t)gk * N + gn) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[kk][ty * TM + i];
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * TN]);
      b[0] = bv.x; b[1] = bv.y; b[2] = bv.z; b[3] = bv.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int gm = m0 + ty * TM + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int gn = n0 + tx * TN + j;
      if (gn >= N) continue;
      float v = acc[i][j] + __ldg(bias + gn);
      if (relu) v = fmaxf(v, 0.f);
      C[(int64_t)gm * N + gn] = v;
    }
  }
}

// In-place row softmax of logits [M, C], C <= 1024.  A row is owned by a group of G lanes
// (G = smallest power of two >= min(C, 32)); max and sum go through __shfl_xor_sync.
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ logits, float* __restrict__ probs, int64_t M, int C,
                    int G) {
  const int lane = threadIdx.x & 31;
  const int sub = lane % G;
  const int rows_per_warp = 32 / G;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r0 = warp * rows_per_warp; r0 < M; r0 += n_warps * rows_per_warp) {
    const int64_t r = r0 + lane / G;
    const bool live = r < M;
    float mx = -INFINITY;
    if (live)
      for (int j = sub; j < C; j += G) mx = fmaxf(mx, logits[r * C + j]);
    for (int o = G >> 1; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    if (live)
      for (int j = sub; j < C; j += G) sum += expf(logits[r * C + j] - mx);
    for (int o = G >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (live)
      for (int j = sub; j < C; j += G) probs[r * C + j] = expf(logits[r * C + j] - mx) / sum;
  }
}

}  // namespace

namespace rfk {

size_t mlp_f32_workspace_bytes(const rfk_mlp* h, int64_t n) {
  // two ping-pong activation buffers of the widest layer output (hidden or logits)
  int widest = h->n_classes;
  for (const Layer& l : h->layers) widest = l.out > widest ? l.out : widest;
  const size_t one = ((size_t)n * widest * sizeof(float) + 255) & ~(size_t)255;
  return 2 * one;
}

int mlp_f32_forward(rfk_mlp* h, const uint8_t* x, int64_t n, float* probs, void* ws, size_t ws_bytes,
                    cudaStream_t st) {
  const size_t need = mlp_f32_workspace_bytes(h, n);
  RFK_REQUIRE(ws != nullptr && ws_bytes >= need, "mlp_forward(f32): workspace too small (%zu < %zu)",
              ws_bytes, need);
  RFK_REQUIRE(n < (int64_t)65535 * BM, "mlp_forward(f32): batch too large for one launch");
  float* buf[2] = {static_cast<float*>(ws), reinterpret_cast<float*>(static_cast<char*>(ws) + need / 2)};
  const void* a = x;
  const int n_layers = (int)h->layers.size();
  for (int i = 0; i < n_layers; ++i) {
    const Layer& l = h->layers[i];
    float* c = buf[i & 1];
    dim3 grid((l.out + BN - 1) / BN, (unsigned)((n + BM - 1) / BM));
    const int relu = i < n_layers - 1;
    if (i == 0)
      sgemm_bias_act_kernel<true><<<grid, kThreads, 0, st>>>(a, l.w_f32, l.b_f32, c, (int)n, l.out, l.in,
                                                            h->scale, h->shift, relu);
    else
      sgemm_bias_act_kernel<false><<<grid, kThreads, 0, st>>>(a, l.w_f32, l.b_f32, c, (int)n, l.out, l.in,
                                                             nullptr, nullptr, relu);
    RFK_LAUNCH_CHECK("sgemm_bias_act_kernel");
    a = c;
  }
  const int C = h->n_classes;
  RFK_REQUIRE(C <= 1024, "mlp_forward(f32): more than 1024 classes");
  int G = 1;
  while (G < C && G < 32) G <<= 1;
  const int64_t rows_per_block = (256 / 32) * (32 / G);
  int64_t grid = (n + rows_per_block - 1) / rows_per_block;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (grid > cap) grid = cap;
  softmax_rows_kernel<<<(int)grid, 256, 0, st>>>(static_cast<const float*>(a), probs, n, C, G);
  RFK_LAUNCH_CHECK("softmax_rows_kernel");
  return RFK_OK;
}

}  // namespace rfk
