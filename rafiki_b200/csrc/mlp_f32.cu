// fp32 parity path of kernel 2: the trial network on CUDA cores, true fp32 FMA.
//
// Replaces the Keras graph of TfFeedForward (TfFeedForward.py:141-157) when the caller
// asks for RFK_F32: results stay within 1e-5 of the fp32 oracle (the tensor-core paths
// round operands to bf16).  Layer 0 reads the uint8 queries directly and applies the
// BatchNormalization affine while staging its A tile, so no decoded copy of the batch is
// written.  The softmax is one warp-slice per row with shuffle reductions.
#include <mutex>

#include "rfk_internal.cuh"

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

// Scratch of the allocation-free entry point rfk_mlp_forward_static (end of this file): a fixed
// block of global memory that is part of the library image on every GPU that loads it.
static __device__ __align__(256) unsigned char g_static_scratch[RFK_STATIC_SCRATCH_BYTES];

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

// ------------------------------------------------------------------------------------------
// rfk_mlp_forward_static: the fp32 trial forward pass for callers that must not allocate.
//
// Some hosts of this library can hand over input and output buffers but cannot obtain scratch
// memory at call time: a serving loop that runs inside a memory budget fixed at start-up, a
// plugin that is given raw device pointers by a framework whose allocator it may not call, a
// process that has handed the rest of the GPU to another tenant.  For them the intermediate
// activations of the fp32 path live in g_static_scratch above instead of a caller workspace.
//
// What a caller should know:
//  * The scratch is RFK_STATIC_SCRATCH_BYTES (4 MiB) of device memory per GPU, held from the
//    moment this part of the library is loaded on that GPU until the process exits.  With CUDA's
//    default lazy module loading that moment can be the FIRST call on the device: the driver then
//    places the module (the 4 MiB block included) on the GPU inside that call, which is an
//    implicit allocation and may synchronise the device.  Later calls allocate nothing.
//  * A batch whose activations do not fit (two ping-pong buffers of n x widest-layer floats,
//    the figure rfk_mlp_workspace_bytes reports) is refused with RFK_ERR_INVALID; split it.
//  * The first call on each device looks up the device address of the scratch block
//    (cudaGetSymbolAddress) with that device current, and remembers it.  Make that first call
//    before starting any stream capture on the device.  The remembered address does not survive
//    cudaDeviceReset: a process that resets a device must not call this entry point on it again.
//  * Every call on a device uses the same block and leaves its activations behind in it.  Nothing
//    in it has to start at zero: each element that a kernel reads was written earlier in the same
//    call.  Two calls that overlap in time on one GPU (different streams, or different threads)
//    would overwrite each other's activations; run them one after the other on one stream.
//  * No cudaMalloc, cudaFree or other allocating routine is called, here or below.
namespace {
std::mutex g_static_mu;
void* g_static_ptr[64] = {nullptr};   // device address of g_static_scratch, per device ordinal
}  // namespace

extern "C" int rfk_mlp_forward_static(rfk_mlp_t* h, const uint8_t* x, int64_t n, float* probs, void* stream) {
  RFK_REQUIRE(h != nullptr, "mlp_forward_static: handle is NULL");
  RFK_REQUIRE(h->dtype == RFK_F32, "mlp_forward_static: only RFK_F32 trials use scratch memory "
              "(the tensor-core path of rfk_mlp_forward needs none and allocates nothing)");
  RFK_REQUIRE(n >= 0, "mlp_forward_static: n=%lld", (long long)n);
  if (n == 0) return RFK_OK;
  RFK_REQUIRE(x != nullptr && probs != nullptr, "mlp_forward_static: NULL buffer");
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  RFK_REQUIRE(dev == h->device, "mlp_forward_static: trial lives on device %d, current device is %d", h->device, dev);
  RFK_REQUIRE(dev >= 0 && dev < 64, "mlp_forward_static: device ordinal %d not supported", dev);
  const size_t need = rfk::mlp_f32_workspace_bytes(h, n);
  if (need > (size_t)RFK_STATIC_SCRATCH_BYTES) {
    rfk::set_error("mlp_forward_static: a batch of %lld queries needs %zu bytes of scratch, the built-in block "
                   "holds %d; use a smaller batch or rfk_mlp_forward with your own workspace",
                   (long long)n, need, RFK_STATIC_SCRATCH_BYTES);
    return RFK_ERR_INVALID;
  }
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (g_static_ptr[dev] == nullptr) {
      // In host code the array's name is only a handle; the kernels need its device address.
      void* p = nullptr;
      RFK_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
      g_static_ptr[dev] = p;
    }
    scratch = g_static_ptr[dev];
  }
  return rfk::mlp_f32_forward(h, x, n, probs, scratch, (size_t)RFK_STATIC_SCRATCH_BYTES,
                              reinterpret_cast<cudaStream_t>(stream));
}
