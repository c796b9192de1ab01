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
#include <stdlib.h>
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

// plain system-scope store: for flag stores that follow ONE __threadfence_system() (a release store each
// would pay a system-wide memory barrier per flag)
__device__ __forceinline__ void st_relaxed_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
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

// Scatter form with SEVERAL trials on this GPU: the trial kernel left per-trial probabilities in the scratch;
// this kernel is kernel 3 and the hand-shake in one -- the rank's partial mean (float64 sum in trial order,
// one division by the size of the whole ensemble, one rounding: ensemble.cu's arithmetic) goes straight into
// the slot of the OWNER of each row with 16-byte stores, after the owners have released the slots
// (`consumed`), and the last block raises `arrived[rank]` on every owner.
struct ScatterMeanArgs {
  const float* trial[RFK_MAX_TRIALS];
  float* owner_out[kMaxRanks];            // this rank's slot in owner j's memory, indexed by global row
  long long owner_vlo[kMaxRanks + 1];     // float4 index where owner j's rows begin
  const unsigned int* wait_flags;
  unsigned int* signal_flags[kMaxRanks];
  unsigned int* counter;
  long long n_vec;
  double divisor;
  unsigned int wait_value, signal_value;
  int w, world;
};
__global__ void __launch_bounds__(kThreads) peer_scatter_mean_kernel(const __grid_constant__ ScatterMeanArgs a) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");     // the next trial kernel may start its prologue
  asm volatile("griddepcontrol.wait;" ::: "memory");                  // launched as a programmatic dependent of the trial kernel
  if (threadIdx.x < a.world) {
    if (!spin_until_ge(a.wait_flags + threadIdx.x, a.wait_value)) {
      printf("rafiki_b200: scatter-mean timed out waiting for owner %d to release its slot\n", (int)threadIdx.x);
      __trap();
    }
  }
  __syncthreads();
  const long long stride = (long long)gridDim.x * kThreads;
  int j = 0;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < a.n_vec; i += stride) {
    while (i >= a.owner_vlo[j + 1]) ++j;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    for (int t = 0; t < a.w; ++t) {
      const float4 v = *reinterpret_cast<const float4*>(a.trial[t] + 4 * i);
      s0 += (double)v.x; s1 += (double)v.y; s2 += (double)v.z; s3 += (double)v.w;
    }
    float4 o;
    o.x = (float)(s0 / a.divisor); o.y = (float)(s1 / a.divisor);
    o.z = (float)(s2 / a.divisor); o.w = (float)(s3 / a.divisor);
    *reinterpret_cast<float4*>(a.owner_out[j] + 4 * i) = o;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(a.counter, 1u) == gridDim.x - 1) {
      atomicExch(a.counter, 0u);
      __threadfence();
      for (int q = 0; q < a.world; ++q) st_relaxed_sys(a.signal_flags[q], a.signal_value);
    }
  }
}

__global__ void peer_wait_kernel(const unsigned int* done_flags, int world, unsigned int step) {
  // it reads nothing but the flags: a kernel queued behind it as a programmatic dependent may start at once
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
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


// =========================================================================================
// One call per serving step (trial-per-GPU ensemble, BASELINE configs[2]).
//
// The host loop of round 1 made three ctypes calls and kept the step counter in Python; at eight
// ranks the Python per step was as long as the kernels.  rfk_peer_t holds the mapped blocks of
// every rank, the layout and the step counters, so that one C call enqueues a whole step:
//   [all-gather of the uint8 query slices over NVLink]  ->  fused trial kernel  ->  cross-rank mean.
// Block layout (the same on every rank, so each rank can compute every peer's addresses):
//   partial[2] | result[2] | queries[2] | slots[3][world] | flags
// where partial / result hold max_n * n_classes floats, queries max_n * d bytes, all rounded up to
// 256 bytes, `slots` (2 x world buffers: where each rank's kernel stores its partial means on the root)
// and the flag words listed in peer_layout().
namespace {

struct PeerLayout {
  size_t buf_bytes, x_bytes, off_partial[2], off_result[2], off_x[2], off_slots[3], off_flags, total;
  int f_arrived, f_consumed, f_counter, f_sumcounter, f_done, f_meancounter;    // indices into the flag words (see below)
};

PeerLayout peer_layout(int world, int64_t max_n, int d, int n_classes) {
  PeerLayout L;
  L.buf_bytes = ((size_t)max_n * (size_t)n_classes * sizeof(float) + 255) / 256 * 256;
  L.x_bytes = ((size_t)max_n * (size_t)d + 255) / 256 * 256;
  size_t off = 0;
  for (int k = 0; k < 2; ++k) { L.off_partial[k] = off; off += L.buf_bytes; }
  for (int k = 0; k < 2; ++k) { L.off_result[k] = off; off += L.buf_bytes; }
  for (int k = 0; k < 2; ++k) { L.off_x[k] = off; off += L.x_bytes; }
  // one slot per source rank, three deep in the scatter form: the kernel of step s overwrites what step s - 3
  // left, which its owner consumed a whole kernel ago, so no rank ever waits for a slower one to free a slot
  for (int k = 0; k < 3; ++k) { L.off_slots[k] = off; off += (size_t)world * L.buf_bytes; }
  L.off_flags = off;
  // flag words: [0,G) partials ready | [G,2G) slice in place | [2G] block counter | [2G+1,3G+1) query slice ready |
  //             [3G+1,4G+1) rank's partial means have arrived in its slot (on every owner) | [4G+1,5G+1) owner j has
  //             consumed its slots (on every rank) | [5G+1] spare | [6G+3,6G+6) block counters of the trial kernel (step % 3) | [5G+2] block counter of
  //             the gather-sum kernel | [5G+3,6G+3) rank's rows of the result are in place (on the root)
  L.f_arrived = 3 * world + 1;
  L.f_consumed = 4 * world + 1;       // G words: owner j has consumed the slots it holds
  L.f_counter = 6 * world + 3;         // three words: the trial kernels of consecutive steps may overlap (step % 3)
  L.f_meancounter = 6 * world + 6;     // block counter of the scatter-mean kernel
  L.f_sumcounter = 5 * world + 2;
  L.f_done = 5 * world + 3;           // G words (root): rank j's rows of the result are in place
  off += ((size_t)(6 * world + 8) * sizeof(unsigned int) + 255) / 256 * 256;
  L.total = off;
  return L;
}

// rows [lo, hi) of the batch that rank r uploads; multiples of 16 rows so that every slice starts
// 16-byte aligned whatever d is
void slice_rows(int64_t n, int world, int r, int64_t* lo, int64_t* hi) {
  const int64_t units = (n + 15) / 16;
  int64_t a = units * r / world * 16, b = units * (r + 1) / world * 16;
  *lo = a < n ? a : n;
  *hi = b < n ? b : n;
}

struct GatherPtrs {
  const uint8_t* x[kMaxRanks];       // query buffer (this step's parity) of every rank, peer mapped
  unsigned int* flags[kMaxRanks];
  long long lo[kMaxRanks], hi[kMaxRanks];   // byte range of every rank's slice
};

// All-gather of the query slices: rank r has uploaded bytes [lo_r, hi_r) of the batch into its own
// buffer (stream-ordered before this kernel); it says so to every peer, waits for theirs, and pulls
// the other G - 1 slices over NVLink into its own buffer with 16-byte loads.
__global__ void __launch_bounds__(kThreads)
peer_allgather_kernel(const GatherPtrs ptrs, uint8_t* __restrict__ mine_x, int world, int rank, unsigned int step) {
  unsigned int* mine = ptrs.flags[rank];
  const int f0 = 2 * world + 1;
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(ptrs.flags[threadIdx.x] + f0 + rank, step);
  }
  if (threadIdx.x < world) {
    if (!spin_until_ge(mine + f0 + threadIdx.x, step)) {
      printf("rafiki_b200: query all-gather timed out waiting for rank %d (rank %d, step %u)\n", (int)threadIdx.x, rank, step);
      __trap();
    }
  }
  __syncthreads();
  const long long stride = (long long)gridDim.x * kThreads * 16;
  for (int jj = 1; jj < world; ++jj) {
    const int j = (rank + jj) % world;             // every rank starts with a different peer
    const uint8_t* src = ptrs.x[j];
    for (long long b = ptrs.lo[j] + ((long long)blockIdx.x * kThreads + threadIdx.x) * 16; b < ptrs.hi[j]; b += 4 * stride) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (b + u * stride < ptrs.hi[j])
          asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
              : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + b + u * stride));
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (b + u * stride < ptrs.hi[j]) *reinterpret_cast<uint4*>(mine_x + b + u * stride) = v[u];
    }
  }
}

// Root side of the gather form: every rank's trial kernel has stored its partial means straight into
// slot[rank] of THIS GPU's block (over NVLink, tile by tile while it computed) and raised arrived[rank].
// One kernel waits for all ranks, adds the slots in float64 in rank order with one rounding (the same sum
// as peer_reduce_kernel), and tells every rank that the slots of this step may be reused.
struct SlotPtrs {
  const float* slot[kMaxRanks];
  unsigned int* consumed[kMaxRanks];    // each rank's "slot consumed" word (peer mapped)
};
__global__ void __launch_bounds__(kThreads)
peer_gather_sum_kernel(const SlotPtrs ptrs, const unsigned int* arrived, int world, unsigned int step, long long v_lo,
                       long long v_hi, float* __restrict__ out, unsigned int* counter, unsigned int* done_flag) {
  if (threadIdx.x < world) {
    if (!spin_until_ge(arrived + threadIdx.x, step)) {
      printf("rafiki_b200: gather-sum timed out waiting for rank %d (step %u)\n", (int)threadIdx.x, step);
      __trap();
    }
  }
  __syncthreads();
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = v_lo + (long long)blockIdx.x * kThreads + threadIdx.x; i < v_hi; i += stride) {
    float4 v[kMaxRanks];
#pragma unroll
    for (int j = 0; j < kMaxRanks; ++j)
      if (j < world) v[j] = __ldcg(reinterpret_cast<const float4*>(ptrs.slot[j]) + i);   // written by peers: bypass L1
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
    for (int j = 0; j < kMaxRanks; ++j)
      if (j < world) {
        s0 += (double)v[j].x; s1 += (double)v[j].y; s2 += (double)v[j].z; s3 += (double)v[j].w;
      }
    float4 o;
    o.x = (float)s0; o.y = (float)s1; o.z = (float)s2; o.w = (float)s3;
    reinterpret_cast<float4*>(out)[i] = o;
  }
  __threadfence_system();      // `out` may be the root's memory
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(counter, 1u) == gridDim.x - 1) {     // last block: every slot has been read, every row stored
      atomicExch(counter, 0u);
      __threadfence_system();
      if (done_flag != nullptr) st_relaxed_sys(done_flag, step);
      for (int j = 0; j < world; ++j) st_relaxed_sys(ptrs.consumed[j], step);
    }
  }
}

}  // namespace

struct rfk_peer {
  int world = 1, rank = 0, root = 0, device = 0, d = 0, n_classes = 0;
  int64_t max_n = 0;
  PeerLayout L;
  uint8_t* base[kMaxRanks] = {nullptr};
  unsigned int rstep = 0, gstep = 0;       // reduce / gather steps issued so far (the same on every rank)
  int grid_blocks = 0;
  // pipelined steps: the step whose slots have not been summed yet (0: none) and its batch size; the result
  // of the newest finished step for callers of the pipelined form
  unsigned int pending_step = 0;
  int64_t pending_n = 0;
  float* held_result = nullptr;
  bool inputs_resident = false;            // rfk_peer_set_inputs_resident
};

extern "C" int rfk_peer_block_bytes(int world, int64_t max_n, int d, int n_classes, size_t* out) {
  RFK_REQUIRE(out != nullptr && world >= 1 && world <= kMaxRanks && max_n >= 1 && d >= 1 && n_classes >= 1,
              "peer_block_bytes: bad argument");
  *out = peer_layout(world, max_n, d, n_classes).total;
  return RFK_OK;
}

extern "C" int rfk_peer_slice_rows(int64_t n, int world, int rank, int64_t* row_lo, int64_t* row_hi) {
  RFK_REQUIRE(row_lo != nullptr && row_hi != nullptr && n >= 0 && world >= 1 && rank >= 0 && rank < world,
              "peer_slice_rows: bad argument");
  slice_rows(n, world, rank, row_lo, row_hi);
  return RFK_OK;
}

extern "C" int rfk_peer_create(rfk_peer_t** out, int world, int rank, int root, void* const* bases, int64_t max_n,
                               int d, int n_classes) {
  RFK_REQUIRE(out != nullptr && bases != nullptr, "peer_create: NULL argument");
  RFK_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world && root >= 0 && root < world,
              "peer_create: bad world/rank/root %d/%d/%d", world, rank, root);
  RFK_REQUIRE(max_n >= 1 && d >= 1 && n_classes >= 1 && (max_n * n_classes) % 4 == 0,
              "peer_create: max_n * n_classes must be a positive multiple of 4");
  rfk_peer* p = new rfk_peer();
  p->world = world; p->rank = rank; p->root = root; p->d = d; p->n_classes = n_classes; p->max_n = max_n;
  p->L = peer_layout(world, max_n, d, n_classes);
  for (int j = 0; j < world; ++j) {
    if (bases[j] == nullptr || (reinterpret_cast<uintptr_t>(bases[j]) & 255) != 0) {
      delete p;
      rfk::set_error("peer_create: block of rank %d is NULL or not 256-byte aligned", j);
      return RFK_ERR_INVALID;
    }
    p->base[j] = static_cast<uint8_t*>(bases[j]);
  }
  if (cudaGetDevice(&p->device) != cudaSuccess) p->device = 0;
  p->grid_blocks = rfk::num_sms();
  *out = p;
  return RFK_OK;
}

extern "C" int rfk_peer_destroy(rfk_peer_t* p) {
  delete p;
  return RFK_OK;
}

extern "C" int rfk_peer_upload_slot(rfk_peer_t* p, int64_t n, uint8_t** dst, int64_t* row_lo, int64_t* row_hi) {
  RFK_REQUIRE(p != nullptr && dst != nullptr && row_lo != nullptr && row_hi != nullptr, "peer_upload_slot: NULL argument");
  RFK_REQUIRE(n >= 1 && n <= p->max_n, "peer_upload_slot: n %lld outside 1..%lld", (long long)n, (long long)p->max_n);
  slice_rows(n, p->world, p->rank, row_lo, row_hi);
  const int k = (int)((p->gstep + 1) & 1);
  *dst = p->base[p->rank] + p->L.off_x[k] + (size_t)*row_lo * (size_t)p->d;
  return RFK_OK;
}

extern "C" int rfk_peer_allgather_u8(rfk_peer_t* p, int64_t n, const uint8_t** gathered, void* stream) {
  RFK_REQUIRE(p != nullptr, "peer_allgather: NULL argument");
  RFK_REQUIRE(n >= 1 && n <= p->max_n, "peer_allgather: n %lld outside 1..%lld", (long long)n, (long long)p->max_n);
  const unsigned int step = ++p->gstep;
  const int k = (int)(step & 1);
  GatherPtrs g;
  for (int j = 0; j < kMaxRanks; ++j) {
    g.x[j] = nullptr; g.flags[j] = nullptr; g.lo[j] = g.hi[j] = 0;
    if (j < p->world) {
      int64_t lo, hi;
      slice_rows(n, p->world, j, &lo, &hi);
      g.x[j] = p->base[j] + p->L.off_x[k];
      g.flags[j] = reinterpret_cast<unsigned int*>(p->base[j] + p->L.off_flags);
      g.lo[j] = lo * p->d;
      g.hi[j] = hi * p->d;
      // slices start at multiples of 16 rows, hence 16-byte aligned; a ragged end is rounded up
      // inside the buffer (x_bytes is padded to 256)
      g.hi[j] = (g.hi[j] + 15) / 16 * 16;
    }
  }
  uint8_t* mine = p->base[p->rank] + p->L.off_x[k];
  peer_allgather_kernel<<<p->grid_blocks, kThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(g, mine, p->world,
                                                                                               p->rank, step);
  RFK_LAUNCH_CHECK("peer_allgather_kernel");
  if (gathered != nullptr) *gathered = mine;
  return RFK_OK;
}

namespace {

// rows (in pair tiles of 256) that rank j owns in the scatter form
inline void owner_bounds(int64_t n, int G, int* bound) {
  const int n_pt = (int)((n + 255) / 256);
  for (int j = 0; j < G; ++j) bound[j] = (int)((long long)n_pt * j / G);
  bound[G] = n_pt;
}

// Scatter form, trial kernel of step `step`: every tile's partial means go to the owner of its rows.  With
// carry_step != 0 the kernel also carries the reduction of that earlier step (batch carry_n) in its idle warp.
int scatter_launch(rfk_peer* p, rfk_mlp_t* const* trials, int w, int total_trials, const uint8_t* x, int64_t n, void* workspace,
                   size_t workspace_bytes, unsigned int step, unsigned int carry_step, int64_t carry_n, bool overlap,
                   void* stream) {
  const int G = p->world, k = (int)(step % 3);
  unsigned int* my_flags = reinterpret_cast<unsigned int*>(p->base[p->rank] + p->L.off_flags);
  unsigned int* root_flags = reinterpret_cast<unsigned int*>(p->base[p->root] + p->L.off_flags);
  rfk::PeerHook hook;
  memset(&hook, 0, sizeof(hook));
  hook.wait_value = step >= 3 ? step - 3 : 0;
  hook.signal_value = step;
  hook.counter = my_flags + p->L.f_counter + k;
  hook.wait_flags = my_flags + p->L.f_consumed;          // one word per owner
  hook.n_wait = G;
  hook.n_signal = G;
  hook.n_owner = G;
  owner_bounds(n, G, hook.owner_bound);
  for (int j = 0; j < G; ++j) {
    hook.signal_flags[j] = reinterpret_cast<unsigned int*>(p->base[j] + p->L.off_flags) + p->L.f_arrived + p->rank;
    hook.out_owner[j] = reinterpret_cast<float*>(p->base[j] + p->L.off_slots[k] + (size_t)p->rank * p->L.buf_bytes);
  }
  if (carry_step != 0) {
    const int kp = (int)(carry_step % 3), kr = (int)(carry_step & 1);
    int bound[kMaxRanks + 1];
    owner_bounds(carry_n, G, bound);
    const long long row_lo = (long long)bound[p->rank] * 256, row_hi_ = (long long)bound[p->rank + 1] * 256;
    const long long row_hi = row_hi_ < carry_n ? row_hi_ : carry_n;
    hook.n_carry = G;
    hook.carry_step = carry_step;
    hook.carry_v_lo = row_lo * p->n_classes / 4;
    hook.carry_v_hi = (row_hi > row_lo ? row_hi : row_lo) * p->n_classes / 4;
    hook.carry_arrived = my_flags + p->L.f_arrived;
    hook.carry_done = root_flags + p->L.f_done + p->rank;
    hook.carry_counter = my_flags + p->L.f_sumcounter;
    hook.carry_out = reinterpret_cast<float*>(p->base[p->root] + p->L.off_result[kr]);
    for (int j = 0; j < G; ++j) {
      hook.carry_slot[j] = reinterpret_cast<const float*>(p->base[p->rank] + p->L.off_slots[kp] + (size_t)j * p->L.buf_bytes);
      hook.carry_consumed[j] = reinterpret_cast<unsigned int*>(p->base[j] + p->L.off_flags) + p->L.f_consumed + p->rank;
    }
  }
  {   // development switches (timing only, results become wrong): RFK_PEER_DEBUG bits 2 no carry, 4 local output, 8 no hand-shake, 16 empty carry
    static const int dbg = []() { const char* e = getenv("RFK_PEER_DEBUG"); return e ? atoi(e) : 0; }();
    if (dbg & 2) hook.n_carry = 0;
    if (dbg & 4) hook.n_owner = 0;
    if (dbg & 8) hook.n_wait = hook.n_signal = 0;
    if (dbg & 16) hook.carry_v_hi = hook.carry_v_lo;      // hand-shake of the carried reduction without its loads and stores
  }
  // overlap with the previous step's kernel only when both fill the chip (then a kernel three steps back, whose
  // block counter this one reuses, has certainly left: every CTA of a step replaces a CTA of the step before)
  hook.no_grid_wait = overlap && w == 1 && (n + 255) / 256 >= rfk::num_sms() / 2 ? 1 : 0;
  if (w == 1) {
    rfk::set_peer_hook(&hook);
    const int rc = rfk_ensemble_forward(trials, 1, total_trials, x, n, hook.out_owner[p->rank], workspace, workspace_bytes, stream);
    if (rfk::peer_hook_pending()) {      // the call did not reach the pair kernel after all
      rfk::set_peer_hook(nullptr);
      rfk::set_error("peer_ensemble_step: internal routing error (hook not consumed)");
      return RFK_ERR_INVALID;
    }
    return rc;
  }
  // ---- several trials on this GPU: the trial kernel carries the previous step's sum and leaves per-trial
  // probabilities in the scratch; the scatter-mean kernel is kernel 3 and the hand-shake in one
  const size_t need = rfk::pair_workspace_bytes(w, p->n_classes, n);
  RFK_REQUIRE(workspace != nullptr && workspace_bytes >= need,
              "peer_ensemble_step: %zu bytes of workspace needed for %d trials per GPU (every rank must pass it: the "
              "ranks choose the form of the step together)", need, w);
  ScatterMeanArgs a;
  memset(&a, 0, sizeof(a));
  for (int t = 0; t < w; ++t)
    a.trial[t] = static_cast<const float*>(workspace) + (size_t)t * (size_t)n * (size_t)p->n_classes;
  for (int j = 0; j < G; ++j) {
    a.owner_out[j] = hook.out_owner[j];
    a.signal_flags[j] = hook.signal_flags[j];
    a.owner_vlo[j] = (long long)hook.owner_bound[j] * 256 * p->n_classes / 4;
  }
  a.n_vec = n * (long long)p->n_classes / 4;
  a.owner_vlo[G] = a.n_vec;
  for (int j = 0; j < G; ++j)
    if (a.owner_vlo[j] > a.n_vec) a.owner_vlo[j] = a.n_vec;
  a.wait_flags = hook.wait_flags;
  a.wait_value = hook.wait_value;
  a.signal_value = hook.signal_value;
  a.counter = my_flags + p->L.f_meancounter;
  a.divisor = (double)total_trials;
  a.w = w;
  a.world = G;
  rfk::PeerHook carry_only = hook;
  carry_only.n_wait = carry_only.n_signal = carry_only.n_owner = 0;
  carry_only.no_grid_wait = 0;
  carry_only.skip_mean = 1;
  rfk::set_peer_hook(&carry_only);
  const int rc = rfk_ensemble_forward(trials, w, total_trials, x, n, hook.out_owner[p->rank], workspace, workspace_bytes, stream);
  if (rfk::peer_hook_pending()) {
    rfk::set_peer_hook(nullptr);
    rfk::set_error("peer_ensemble_step: internal routing error (hook not consumed)");
    return RFK_ERR_INVALID;
  }
  if (rc != RFK_OK) return rc;
  const long long want = (a.n_vec + kThreads - 1) / kThreads;
  const int grid = (int)(want < 1 ? 1 : (want < (long long)rfk::num_sms() * 4 ? want : (long long)rfk::num_sms() * 4));
  {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = getenv("RFK_NO_PDL") == nullptr ? 1 : 0;
    RFK_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_scatter_mean_kernel, a));
  }
  RFK_LAUNCH_CHECK("peer_scatter_mean_kernel");
  return RFK_OK;
}

// Scatter form, stand-alone reduction of step `step`: my rows of the G slots -> the root's result; the root then
// waits for every rank's rows.
int scatter_sum(rfk_peer* p, unsigned int step, int64_t n, void* stream) {
  const int G = p->world, k = (int)(step % 3), kr = (int)(step & 1);
  unsigned int* my_flags = reinterpret_cast<unsigned int*>(p->base[p->rank] + p->L.off_flags);
  unsigned int* root_flags = reinterpret_cast<unsigned int*>(p->base[p->root] + p->L.off_flags);
  int bound[kMaxRanks + 1];
  owner_bounds(n, G, bound);
  const long long row_lo = (long long)bound[p->rank] * 256, row_hi_ = (long long)bound[p->rank + 1] * 256;
  const long long row_hi = row_hi_ < n ? row_hi_ : n;
  SlotPtrs sp;
  for (int j = 0; j < kMaxRanks; ++j) {
    sp.slot[j] = nullptr;
    sp.consumed[j] = nullptr;
    if (j < G) {
      sp.slot[j] = reinterpret_cast<const float*>(p->base[p->rank] + p->L.off_slots[k] + (size_t)j * p->L.buf_bytes);
      sp.consumed[j] = reinterpret_cast<unsigned int*>(p->base[j] + p->L.off_flags) + p->L.f_consumed + p->rank;
    }
  }
  float* res = reinterpret_cast<float*>(p->base[p->root] + p->L.off_result[kr]);
  const long long v_lo = row_lo * p->n_classes / 4, v_hi = (row_hi > row_lo ? row_hi : row_lo) * p->n_classes / 4;
  const long long want = (v_hi - v_lo + kThreads - 1) / kThreads;
  const int grid = (int)(want < 1 ? 1 : (want < (long long)p->grid_blocks ? want : (long long)p->grid_blocks));
  peer_gather_sum_kernel<<<grid, kThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      sp, my_flags + p->L.f_arrived, G, step, v_lo, v_hi, res, my_flags + p->L.f_sumcounter, root_flags + p->L.f_done + p->rank);
  RFK_LAUNCH_CHECK("peer_gather_sum_kernel");
  if (p->rank == p->root) return rfk_peer_wait_flags(root_flags + p->L.f_done, G, step, stream);
  return RFK_OK;
}

// Every rank must reach the same answer: with trials dealt t % world a rank holds at least one trial exactly
// when total_trials >= world, and the other conditions do not depend on the rank.
bool scatter_capable(const rfk_peer* p, rfk_mlp_t* const* trials, int w, int total_trials, int64_t n) {
  if (w < 1 || total_trials < p->world || p->world <= 1 || trials == nullptr) return false;
  static const bool multi = getenv("RFK_PEER_NO_MULTI_SCATTER") == nullptr;
  if (total_trials != p->world && !multi) return false;      // (a rank-independent condition)
  for (int t = 0; t < w; ++t)
    if (trials[t] == nullptr || !trials[t]->fusable || trials[t]->dtype == RFK_F32 || trials[t]->dtype != trials[0]->dtype ||
        trials[t]->d != trials[0]->d)
      return false;
  return rfk::pair_applicable(trials[0]->d, n);
}

}  // namespace

extern "C" int rfk_peer_ensemble_step(rfk_peer_t* p, rfk_mlp_t* const* trials, int w, int total_trials,
                                      const uint8_t* x, int64_t n, void* workspace, size_t workspace_bytes,
                                      float** result, void* stream) {
  RFK_REQUIRE(p != nullptr, "peer_ensemble_step: NULL argument");
  RFK_REQUIRE(n >= 1 && n <= p->max_n && (n * p->n_classes) % 4 == 0,
              "peer_ensemble_step: n %lld outside 1..%lld or n * classes not a multiple of 4", (long long)n,
              (long long)p->max_n);
  RFK_REQUIRE(x != nullptr, "peer_ensemble_step: x is NULL (pass the pointer rfk_peer_allgather_u8 returned)");
  if (p->pending_step != 0) {          // a pipelined step is still open: finish it first
    const int rc0 = scatter_sum(p, p->pending_step, p->pending_n, stream);
    if (rc0 != RFK_OK) return rc0;
    p->pending_step = 0;
  }
  const unsigned int step = ++p->rstep;
  const int k = (int)(step & 1);
  const int64_t count = n * p->n_classes;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  static const bool reduce_mode = []() { const char* m = getenv("RFK_PEER_MODE"); return m != nullptr && m[0] == 'r'; }();
  if (!reduce_mode) {
    // Gather forms (default): a rank's trial kernel stores its partial means straight into slots in OTHER
    // GPUs' blocks while it computes (16-byte stores over NVLink from its epilogue), so no communication
    // kernel stands between two trial kernels.
    //  * scatter (one trial per rank, batch large enough for the pair kernel): the rows of the batch are
    //    divided among the ranks; every tile goes to the OWNER of its rows; each owner adds the G slots of
    //    its rows (float64, rank order, one rounding) and stores the result rows into the root's buffer.
    //    Every GPU does 1/G of the summing and no GPU waits for a slower one inside its trial kernel.
    //  * fixed root (every other case): all slots live on the root, which adds them.
    // Hand-shake per step s, slots double buffered by parity: a slot may be overwritten once its owner has
    // consumed step s - 2; after its last store a rank raises arrived[rank] on every owner.
    unsigned int* my_flags = reinterpret_cast<unsigned int*>(p->base[p->rank] + p->L.off_flags);
    unsigned int* root_flags = reinterpret_cast<unsigned int*>(p->base[p->root] + p->L.off_flags);
    const int G = p->world;
    const bool scatter = scatter_capable(p, trials, w, total_trials, n);
    const unsigned int wait_value = step >= 2 ? step - 2 : 0;
    if (scatter) {
      int rc = scatter_launch(p, trials, w, total_trials, x, n, workspace, workspace_bytes, step, 0, 0, false, stream);
      if (rc != RFK_OK) return rc;
      rc = scatter_sum(p, step, n, stream);
      if (rc != RFK_OK) return rc;
      if (result != nullptr) *result = reinterpret_cast<float*>(p->base[p->rank] + p->L.off_result[k]);
      return RFK_OK;
    }
    // ---- fixed root
    float* slot = reinterpret_cast<float*>(p->base[p->root] + p->L.off_slots[k] + (size_t)p->rank * p->L.buf_bytes);
    const unsigned int* wait_flag = my_flags + p->L.f_consumed + p->root;
    unsigned int* signal_flag = root_flags + p->L.f_arrived + p->rank;
    if (wait_value > 0) {
      const int rc = rfk_peer_wait_flags(wait_flag, 1, wait_value, stream);
      if (rc != RFK_OK) return rc;
    }
    if (w > 0) {
      RFK_REQUIRE(trials != nullptr, "peer_ensemble_step: trials is NULL");
      const int rc = rfk_ensemble_forward(trials, w, total_trials, x, n, slot, workspace, workspace_bytes, stream);
      if (rc != RFK_OK) return rc;
    } else {      // a rank that holds no trial (fewer trials than GPUs) contributes zeros
      RFK_CUDA_TRY(cudaMemsetAsync(slot, 0, (size_t)count * sizeof(float), st));
    }
    {
      unsigned int* sig[1] = {signal_flag};
      const int rc = rfk_peer_signal(sig, 1, step, stream);
      if (rc != RFK_OK) return rc;
    }
    if (p->rank == p->root) {
      SlotPtrs sp;
      for (int j = 0; j < kMaxRanks; ++j) {
        sp.slot[j] = nullptr;
        sp.consumed[j] = nullptr;
        if (j < G) {
          sp.slot[j] = reinterpret_cast<const float*>(p->base[p->root] + p->L.off_slots[k] + (size_t)j * p->L.buf_bytes);
          sp.consumed[j] = reinterpret_cast<unsigned int*>(p->base[j] + p->L.off_flags) + p->L.f_consumed + p->root;
        }
      }
      float* res = reinterpret_cast<float*>(p->base[p->root] + p->L.off_result[k]);
      const long long n_vec = count / 4;
      const long long want = (n_vec + kThreads - 1) / kThreads;
      const int grid = (int)(want < (long long)p->grid_blocks * 4 ? want : (long long)p->grid_blocks * 4);
      peer_gather_sum_kernel<<<grid, kThreads, 0, st>>>(sp, root_flags + p->L.f_arrived, G, step, 0, n_vec, res,
                                                        root_flags + p->L.f_sumcounter, nullptr);
      RFK_LAUNCH_CHECK("peer_gather_sum_kernel");
    }
    if (result != nullptr) *result = reinterpret_cast<float*>(p->base[p->rank] + p->L.off_result[k]);
    return RFK_OK;
  }
  // Reduce-scatter form (RFK_PEER_MODE=reduce): partial means in local memory, every rank sums 1/G of the rows
  float* partial = reinterpret_cast<float*>(p->base[p->rank] + p->L.off_partial[k]);
  if (w > 0) {
    RFK_REQUIRE(trials != nullptr, "peer_ensemble_step: trials is NULL");
    const int rc = rfk_ensemble_forward(trials, w, total_trials, x, n, partial, workspace, workspace_bytes, stream);
    if (rc != RFK_OK) return rc;
  } else {      // a rank that holds no trial (fewer trials than GPUs) contributes zeros
    RFK_CUDA_TRY(cudaMemsetAsync(partial, 0, (size_t)count * sizeof(float), st));
  }
  const float* probs[kMaxRanks];
  unsigned int* flags[kMaxRanks];
  for (int j = 0; j < p->world; ++j) {
    probs[j] = reinterpret_cast<const float*>(p->base[j] + p->L.off_partial[k]);
    flags[j] = reinterpret_cast<unsigned int*>(p->base[j] + p->L.off_flags);
  }
  float* root_out = reinterpret_cast<float*>(p->base[p->root] + p->L.off_result[k]);
  int rc = rfk_peer_reduce(probs, flags, p->world, p->rank, p->root, step, count, root_out, p->grid_blocks, stream);
  if (rc != RFK_OK) return rc;
  if (p->rank == p->root) {
    rc = rfk_peer_wait(flags[p->rank], p->world, step, stream);
    if (rc != RFK_OK) return rc;
  }
  if (result != nullptr) *result = reinterpret_cast<float*>(p->base[p->rank] + p->L.off_result[k]);
  return RFK_OK;
}


// Pipelined form for back-to-back batches.  Same work as rfk_peer_ensemble_step, but in the scatter form the
// reduction of step s rides inside the trial kernel of step s + 1 (an idle warp of every CTA adds this GPU's
// rows while the tensor cores work), so a step is ONE kernel per rank.  *prev_result receives the result buffer
// of the PREVIOUS step (NULL after the first call); on the root it is complete once `stream` has passed this
// call.  rfk_peer_flush finishes the last step.  Configurations the scatter form does not take run the plain
// step and hand its result out one call later, so callers see one behaviour.
extern "C" int rfk_peer_ensemble_step_pipelined(rfk_peer_t* p, rfk_mlp_t* const* trials, int w, int total_trials,
                                                const uint8_t* x, int64_t n, void* workspace, size_t workspace_bytes,
                                                float** prev_result, void* stream) {
  RFK_REQUIRE(p != nullptr, "peer_ensemble_step_pipelined: NULL argument");
  static const bool reduce_mode = []() { const char* m = getenv("RFK_PEER_MODE"); return m != nullptr && m[0] == 'r'; }();
  if (prev_result != nullptr) *prev_result = nullptr;
  if (reduce_mode || !scatter_capable(p, trials, w, total_trials, n)) {
    float* prev = p->held_result;
    if (p->pending_step != 0) {
      const int rc0 = rfk_peer_flush(p, &prev, stream);
      if (rc0 != RFK_OK) return rc0;
    }
    float* cur = nullptr;
    const int rc = rfk_peer_ensemble_step(p, trials, w, total_trials, x, n, workspace, workspace_bytes, &cur, stream);
    if (rc != RFK_OK) return rc;
    p->held_result = cur;
    if (prev_result != nullptr) *prev_result = prev;
    return RFK_OK;
  }
  RFK_REQUIRE(n >= 1 && n <= p->max_n && (n * p->n_classes) % 4 == 0,
              "peer_ensemble_step_pipelined: n %lld outside 1..%lld or n * classes not a multiple of 4", (long long)n,
              (long long)p->max_n);
  RFK_REQUIRE(x != nullptr, "peer_ensemble_step_pipelined: x is NULL");
  float* prev = p->held_result;
  p->held_result = nullptr;
  const unsigned int step = ++p->rstep;
  const unsigned int carry = p->pending_step;
  const int rc = scatter_launch(p, trials, w, total_trials, x, n, workspace, workspace_bytes, step, carry, p->pending_n,
                                p->inputs_resident, stream);
  if (rc != RFK_OK) return rc;
  if (carry != 0) {
    static const int dbg = []() { const char* e = getenv("RFK_PEER_DEBUG"); return e ? atoi(e) : 0; }();
    if (p->rank == p->root && !(dbg & 1)) {      // every rank's rows of the carried step are in place before the caller reads them
      unsigned int* root_flags = reinterpret_cast<unsigned int*>(p->base[p->root] + p->L.off_flags);
      if (p->inputs_resident) {
        // queued as a programmatic dependent: the stream still does not pass it before the rows are in place, but
        // the next step's trial kernel need not wait for it to END before it starts
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(1);
        cfg.blockDim = dim3(32);
        cfg.stream = reinterpret_cast<cudaStream_t>(stream);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        RFK_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_wait_kernel, (const unsigned int*)(root_flags + p->L.f_done), p->world, carry));
        RFK_LAUNCH_CHECK("peer_wait_kernel");
      } else {
        const int rc2 = rfk_peer_wait_flags(root_flags + p->L.f_done, p->world, carry, stream);
        if (rc2 != RFK_OK) return rc2;
      }
    }
    prev = reinterpret_cast<float*>(p->base[p->rank] + p->L.off_result[carry & 1]);
  }
  p->pending_step = step;
  p->pending_n = n;
  if (prev_result != nullptr) *prev_result = prev;
  return RFK_OK;
}

extern "C" int rfk_peer_set_inputs_resident(rfk_peer_t* p, int on) {
  RFK_REQUIRE(p != nullptr, "peer_set_inputs_resident: NULL argument");
  p->inputs_resident = on != 0;
  return RFK_OK;
}

extern "C" int rfk_peer_flush(rfk_peer_t* p, float** result, void* stream) {
  RFK_REQUIRE(p != nullptr, "peer_flush: NULL argument");
  float* res = p->held_result;
  p->held_result = nullptr;
  if (p->pending_step != 0) {
    const int rc = scatter_sum(p, p->pending_step, p->pending_n, stream);
    if (rc != RFK_OK) return rc;
    res = reinterpret_cast<float*>(p->base[p->rank] + p->L.off_result[p->pending_step & 1]);
    p->pending_step = 0;
  }
  if (result != nullptr) *result = res;
  return RFK_OK;
}
