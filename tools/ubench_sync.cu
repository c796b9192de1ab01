// Development micro-benchmark: mbarrier hand-off latency between warps and tcgen05 commit latency.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I rafiki_b200/csrc tools/ubench_sync.cu -o tools/build/ubench_sync
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "sm100.cuh"
using namespace sm100;

__device__ __forceinline__ void wait_spin(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void wait_test(uint64_t* bar, uint32_t parity) {
  while (!mbar_test(bar, parity)) {
  }
}

// mode 0: ping-pong, hint wait (library mbar_wait); 1: ping-pong, try_wait spin; 2: ping-pong, test_wait spin
// mode 3: n MMAs (N=128, SS) + commit, same warp waits (hint); 4: same, spin; 5: commit only (no MMA), hint
// mode 6: like 3 but a DIFFERENT warp waits and replies through a second barrier (full hand-off loop)
__global__ void __launch_bounds__(128, 1) ubench(int mode, int iters, int n_mma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  long long t = 0;
  if (mode <= 2) {
    if (warp == 0) {
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        if (lane == 0) mbar_arrive(&bar[0]);
        __syncwarp();
        if (mode == 0) mbar_wait(&bar[1], i & 1); else if (mode == 1) wait_spin(&bar[1], i & 1); else wait_test(&bar[1], i & 1);
      }
      t = clock64() - t0;
    } else if (warp == 1) {
      for (int i = 0; i < iters; ++i) {
        if (mode == 0) mbar_wait(&bar[0], i & 1); else if (mode == 1) wait_spin(&bar[0], i & 1); else wait_test(&bar[0], i & 1);
        if (lane == 0) mbar_arrive(&bar[1]);
        __syncwarp();
      }
    }
  } else if (mode <= 5) {
    if (warp == 0) {
      const uint32_t idesc = idesc_16bit_f32(true, 128, 128);
      const uint32_t a_lo = desc_lo_sw128(smem_u32(smem)), b_lo = desc_lo_sw128(smem_u32(smem) + 16384);
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        if (mode != 5)
          for (int k = 0; k < n_mma; k += 4) mma4_f16_ss_elect(0u, desc_from_lo(a_lo), desc_from_lo(b_lo), idesc, 1u, 4);
        mma_commit_elect(&bar[0]);
        if (mode == 4) wait_spin(&bar[0], i & 1); else mbar_wait(&bar[0], i & 1);
        tc_fence_after_sync();
      }
      t = clock64() - t0;
    }
  } else if (mode == 7) {
    // probe of a phase that completed long ago, result feeds the next probe's address (serial chain)
    if (warp == 0) {
      if (lane == 0) mbar_arrive(&bar[0]);
      __syncwarp();
      uint32_t acc = 0;
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) acc += mbar_try_wait(&bar[acc & 0u], 0u) ? 1u : 0u;
      t = clock64() - t0 + (acc == 12345u ? 1 : 0);
    }
  } else if (mode == 8 || mode == 9 || mode == 10) {
    // issue rate: n_mma MMAs + one commit per iteration, nobody waits until the end (bar[0] counts phases freely)
    if (warp == 0) {
      const uint32_t idesc = idesc_16bit_f32(true, 128, mode == 8 ? 128 : 16);    // 9, 10: tiny MMAs, the issue path shows
      const uint32_t a_lo = desc_lo_sw128(smem_u32(smem)), b_lo = desc_lo_sw128(smem_u32(smem) + 16384);
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        for (int k = 0; k < n_mma; k += 4) mma4_f16_ss_elect(0u, desc_from_lo(a_lo), desc_from_lo(b_lo), idesc, 1u, 4);
        if (mode != 10) mma_commit_elect(&bar[0]);
      }
      mma_commit_elect(&bar[1]);
      mbar_wait(&bar[1], 0);
      t = clock64() - t0;
    }
  } else {
    if (warp == 0) {
      const uint32_t idesc = idesc_16bit_f32(true, 128, 128);
      const uint32_t a_lo = desc_lo_sw128(smem_u32(smem)), b_lo = desc_lo_sw128(smem_u32(smem) + 16384);
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        for (int k = 0; k < n_mma; k += 4) mma4_f16_ss_elect(0u, desc_from_lo(a_lo), desc_from_lo(b_lo), idesc, 1u, 4);
        mma_commit_elect(&bar[0]);
        mbar_wait(&bar[1], i & 1);
        tc_fence_after_sync();
      }
      t = clock64() - t0;
    } else if (warp == 1) {
      for (int i = 0; i < iters; ++i) {
        mbar_wait(&bar[0], i & 1);
        tc_fence_after_sync();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar[1]);
      }
    }
  }
  if (threadIdx.x == 0) out[blockIdx.x] = t;
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(0u);
}

int main() {
  long long* d;
  const int grid = 148;
  cudaMalloc(&d, grid * sizeof(long long));
  cudaFuncSetAttribute(ubench, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * 1024);
  std::vector<long long> h(grid);
  auto run = [&](const char* name, int mode, int iters, int n_mma) {
    ubench<<<grid, 128, 32 * 1024>>>(mode, iters, n_mma, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); exit(1); }
    cudaMemcpy(h.data(), d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    double s = 0; for (auto v : h) s += v; s /= grid;
    printf("%-64s %8.1f cycles per iteration\n", name, s / iters);
  };
  run("warmup", 0, 100, 0);
  run("ping-pong (2 hand-offs), try_wait + suspend hint", 0, 2000, 0);
  run("ping-pong (2 hand-offs), try_wait spin", 1, 2000, 0);
  run("ping-pong (2 hand-offs), test_wait spin", 2, 2000, 0);
  run("try_wait on a completed phase, serial chain", 7, 2000, 0);
  run("issue only: 4 MMAs (N128) + commit per iteration, no waits", 8, 2000, 4);
  run("issue only: 8 MMAs (N128) + commit per iteration, no waits", 8, 2000, 8);
  run("issue only: commit per iteration, no MMAs", 8, 2000, 0);
  run("issue only: 4 MMAs (N16) + commit per iteration", 9, 2000, 4);
  run("issue only: 4 MMAs (N16), no commit", 10, 2000, 4);
  run("issue only: 16 MMAs (N16), no commit", 10, 2000, 16);
  run("commit only + wait (hint)", 5, 2000, 0);
  for (int n : {4, 8, 16}) {
    char buf[96];
    snprintf(buf, sizeof buf, "%d MMAs (N128) + commit + wait, same warp, hint", n); run(buf, 3, 1000, n);
    snprintf(buf, sizeof buf, "%d MMAs (N128) + commit + wait, same warp, spin", n); run(buf, 4, 1000, n);
    snprintf(buf, sizeof buf, "%d MMAs + commit -> other warp waits -> arrive -> wait", n); run(buf, 6, 1000, n);
  }
  return 0;
}
