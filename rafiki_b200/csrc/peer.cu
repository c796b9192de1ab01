// Kernel 3 across GPUs: the cross-trial mean fused with the collective, over NVLink peer memory.
//
// Replaces the Predictor's gather of one answer per worker and ensemble_predictions
// (rafiki/predictor/predictor.py:46-70, rafiki/predictor/ensemble.py:10-14) when the workers are
// one-process-per-GPU ranks of one box.  Each rank's fused kernel has left its partial mean
// (sum over its trials of p_t / W) in a buffer that every other rank has mapped (CUDA IPC).  This
// kernel, launched on every rank right after it:
//   1. tells every peer "my partials of step s are ready" (one system-scope store per peer),
//   2. waits until all peers have said so,
//   3. reads ITS OWN 1/G slice of the rows from all G buffers over NVLink (a reduce-scatter:
//      every link carries 1/G of the data), sums in float64 in rank order, rounds once, and
//      stores the slice straight into the root rank's result buffer (the gather),
//   4. tells the root "my slice of step s is in place".
// rfk_peer_wait (root only) then holds the stream until all slices have arrived.
// NCCL's reduce of the same 2.6 MB costs ~68 us per step at 8 ranks (latency-bound); here every
// GPU moves 1/G of the bytes with plain 16-byte loads.
//
// Flags are monotonic step counters, so nothing is ever reset; buffers are double buffered by
// step parity by the caller (rafiki_b200/distributed.py).  Every spin is bounded: a rank that
// died makes the others trap after ~2 s instead of hanging the box.
#include <string.h>

#include "rfk_internal.cuh"

namespace {

// Measured on 8 GPUs: a register-light variant of this kernel that can co-reside with the next
// batch's fused kernel (which holds 640 x 96 registers per SM) was slower than letting this one run
// alone between two fused kernels, so it is written for its own speed: 256 threads, every rank's
// load issued before the first add.
constexpr int kThreads = 256;
constexpr int kMaxRanks = 16;

struct PeerPtrs {
  const float* probs[kMaxRanks];     // partial means of every rank (peer mapped), this step's buffer
  unsigned int* flags[kMaxRanks];    // flag block of every rank (peer mapped): [0,G) ready, [G,2G) done, [2G] counter
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool spin_until_ge(const unsigned int* p, unsigned int want) {
  const long long t0 = clock64();
  while ((int)(ld_acquire_sys(p) - want) < 0) {
    if (clock64() - t0 > 4000000000LL) return false;
  }
  return true;
}
// The partials were complete before their owner raised its flag (acquired above) and do not change
// while this kernel runs, so the loads may be hoisted and overlapped freely (non-coherent path;
// the L1 is invalidated at every kernel launch, so nothing stale survives from the step before).
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 r;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

__global__ void __launch_bounds__(kThreads)
peer_reduce_kernel(const PeerPtrs ptrs, int world, int rank, int root, unsigned int step, long long n_vec,
                   float* __restrict__ root_out) {
  unsigned int* mine = ptrs.flags[rank];
  // 1. publish "ready": the fused kernel that wrote my partials precedes this kernel on the stream
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(ptrs.flags[threadIdx.x] + rank, step);
  }
  // 2. wait for every rank's "ready"
  if (threadIdx.x < world) {
    if (!spin_until_ge(mine + threadIdx.x, step)) {
      printf("rafiki_b200: peer reduce timed out waiting for rank %d (rank %d, step %u)\n", (int)threadIdx.x, rank, step);
      __trap();
    }
  }
  __syncthreads();
  // 3. my slice of the rows: float64 sum in rank order, one rounding, store into the root's buffer
  const long long lo = n_vec * rank / world, hi = n_vec * (rank + 1) / world;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = lo + (long long)blockIdx.x * kThreads + threadIdx.x; i < hi; i += stride) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    // all ranks' loads are issued before the first sum: NVLink latency is paid once per element
    float4 v[kMaxRanks];
#pragma unroll
    for (int j = 0; j < kMaxRanks; ++j)
      if (j < world) v[j] = ld_peer_f4(ptrs.probs[j] + 4 * i);
#pragma unroll
    for (int j = 0; j < kMaxRanks; ++j)
      if (j < world) {
        s0 += (double)v[j].x; s1 += (double)v[j].y; s2 += (double)v[j].z; s3 += (double)v[j].w;
      }
    float4 o;
    o.x = (float)s0; o.y = (float)s1; o.z = (float)s2; o.w = (float)s3;
    *reinterpret_cast<float4*>(root_out + 4 * i) = o;
  }
  // 4. last block out tells the root that this rank's slice is in place
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* counter = mine + 2 * world;
    const unsigned int done = atomicAdd(counter, 1u) + 1u;
    if (done == step * gridDim.x) {          // the counter is never reset: step-th multiple of the grid size
      __threadfence_system();
      st_release_sys(ptrs.flags[root] + world + rank, step);
    }
  }
}

struct FlagPtrs {
  unsigned int* p[kMaxRanks];
};
__global__ void peer_signal_kernel(const FlagPtrs flags, int count, unsigned int step) {
  // everything this stream did before (the fused kernel's stores into peer memory) is ordered first
  __threadfence_system();
  if (threadIdx.x < count) st_release_sys(flags.p[threadIdx.x], step);
}

__global__ void peer_wait_kernel(const unsigned int* done_flags, int world, unsigned int step) {
  if (threadIdx.x < world) {
    if (!spin_until_ge(done_flags + threadIdx.x, step)) {
      printf("rafiki_b200: peer wait timed out on rank %d (step %u)\n", (int)threadIdx.x, step);
      __trap();
    }
  }
}

}  // namespace

// One device allocation that other processes on the box can map: cudaMalloc + IPC handle (64 bytes).
// Set-up-time calls (they allocate and synchronise); the data path above allocates nothing.
extern "C" int rfk_peer_alloc(size_t bytes, void** dptr, void* handle64) {
  RFK_REQUIRE(bytes > 0 && dptr != nullptr && handle64 != nullptr, "peer_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  RFK_CUDA_TRY(cudaMalloc(&p, bytes));
  RFK_CUDA_TRY(cudaMemset(p, 0, bytes));
  RFK_CUDA_TRY(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  const cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    RFK_CUDA_TRY(e);
  }
  memcpy(handle64, &h, 64);
  *dptr = p;
  return RFK_OK;
}
// Map another process's rfk_peer_alloc block into this process, for kernels of the CURRENT device.
extern "C" int rfk_peer_open(const void* handle64, void** dptr) {
  RFK_REQUIRE(handle64 != nullptr && dptr != nullptr, "peer_open: bad argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  RFK_CUDA_TRY(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RFK_OK;
}
extern "C" int rfk_peer_close(void* dptr) {
  if (dptr != nullptr) RFK_CUDA_TRY(cudaIpcCloseMemHandle(dptr));
  return RFK_OK;
}
extern "C" int rfk_peer_free(void* dptr) {
  if (dptr != nullptr) RFK_CUDA_TRY(cudaFree(dptr));
  return RFK_OK;
}

extern "C" int rfk_peer_enable(int peer_device) {
  int dev = 0;
  RFK_CUDA_TRY(cudaGetDevice(&dev));
  if (peer_device == dev) return RFK_OK;
  int can = 0;
  RFK_CUDA_TRY(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  if (!can) {
    rfk::set_error("peer_enable: device %d cannot map the memory of device %d (no NVLink / P2P path)", dev, peer_device);
    return RFK_ERR_COMM;
  }
  const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();   // clear the sticky-free error state
    return RFK_OK;
  }
  RFK_CUDA_TRY(e);
  return RFK_OK;
}

extern "C" int rfk_peer_reduce(const float* const* probs, unsigned int* const* flags, int world, int rank, int root,
                               unsigned int step, int64_t count, float* root_out, int grid_blocks, void* stream) {
  RFK_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world && root >= 0 && root < world,
              "peer_reduce: bad world/rank/root %d/%d/%d", world, rank, root);
  RFK_REQUIRE(probs != nullptr && flags != nullptr && root_out != nullptr, "peer_reduce: NULL argument");
  RFK_REQUIRE(step >= 1, "peer_reduce: steps count from 1");
  RFK_REQUIRE(count > 0 && count % 4 == 0, "peer_reduce: count must be a positive multiple of 4");
  RFK_REQUIRE(grid_blocks >= 1, "peer_reduce: grid_blocks must be the same positive number on every call");
  PeerPtrs p;
  for (int j = 0; j < kMaxRanks; ++j) {
    p.probs[j] = j < world ? probs[j] : nullptr;
    p.flags[j] = j < world ? flags[j] : nullptr;
    if (j < world) {
      RFK_REQUIRE(probs[j] != nullptr && flags[j] != nullptr, "peer_reduce: rank %d pointers are NULL", j);
      RFK_REQUIRE((reinterpret_cast<uintptr_t>(probs[j]) & 15) == 0, "peer_reduce: buffers must be 16-byte aligned");
    }
  }
  RFK_REQUIRE((reinterpret_cast<uintptr_t>(root_out) & 15) == 0, "peer_reduce: result buffer must be 16-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  peer_reduce_kernel<<<grid_blocks, kThreads, 0, st>>>(p, world, rank, root, step, count / 4, root_out);
  RFK_LAUNCH_CHECK("peer_reduce_kernel");
  return RFK_OK;
}

extern "C" int rfk_peer_signal(unsigned int* const* flags, int count, unsigned int step, void* stream) {
  RFK_REQUIRE(flags != nullptr && count >= 1 && count <= kMaxRanks, "peer_signal: bad argument");
  FlagPtrs f;
  for (int i = 0; i < kMaxRanks; ++i) f.p[i] = i < count ? flags[i] : nullptr;
  for (int i = 0; i < count; ++i) RFK_REQUIRE(flags[i] != nullptr, "peer_signal: flag %d is NULL", i);
  peer_signal_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(f, count, step);
  RFK_LAUNCH_CHECK("peer_signal_kernel");
  return RFK_OK;
}

extern "C" int rfk_peer_wait_flags(const unsigned int* flags, int count, unsigned int step, void* stream) {
  RFK_REQUIRE(flags != nullptr && count >= 1 && count <= kMaxRanks, "peer_wait_flags: bad argument");
  peer_wait_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flags, count, step);
  RFK_LAUNCH_CHECK("peer_wait_kernel");
  return RFK_OK;
}

extern "C" int rfk_peer_wait(const unsigned int* flags_of_this_rank, int world, unsigned int step, void* stream) {
  RFK_REQUIRE(flags_of_this_rank != nullptr && world >= 1 && world <= kMaxRanks, "peer_wait: bad argument");
  peer_wait_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flags_of_this_rank + world, world, step);
  RFK_LAUNCH_CHECK("peer_wait_kernel");
  return RFK_OK;
}
