// Development micro-benchmark (not part of the library): tensor-memory load/store and tcgen05.mma
// issue rates on one SM, measured with clock64 inside the kernel.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I rafiki_b200/csrc tools/ubench_tmem.cu -o tools/build/ubench_tmem
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "sm100.cuh"

using namespace sm100;

__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&v)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
      "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]), "=r"(v[32]),
        "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]),
        "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]),
        "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]),
        "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
      "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}

// mode: 0 ld x16 + wait per round; 1 four ld x16 then one wait; 2 ld x32 + wait; 3 ld x64 + wait;
//       4 st x32 + wait::st; 5 eight st x8 + one wait::st
// n_ld_warps: warps doing the TMEM traffic (warp w touches lane quarter w % 4)
// mma_mode: 0 none; 1 SS N=128; 2 SS N=256; 3 TS N=128; 4 TS N=256  (issued by warp 16, endlessly until the others finish)
struct Result {
  long long ld_cycles;     // per CTA: cycles for `rounds` rounds (warp 0)
  long long mma_cycles;    // cycles for n_mma MMAs, issue to completion
  long long mma_issue;     // cycles until the last MMA was issued
  unsigned sink;
};

__global__ void __launch_bounds__(576, 1)
ubench(int mode, int n_ld_warps, int rounds, int mma_mode, int n_mma, Result* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar_done;
  __shared__ uint32_t tmem_slot;
  __shared__ volatile int ld_done;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // 1.0h
  if (threadIdx.x == 0) {
    mbar_init(&bar_done, 1);
    fence_mbar_init();
    ld_done = 0;
  }
  if (warp == 16) tmem_alloc<512>(&tmem_slot);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  unsigned sink = 0;
  long long t_ld = 0, t_mma = 0, t_issue = 0;
  if (warp < n_ld_warps) {
    const uint32_t lane_addr = ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t col0 = (uint32_t)((warp >> 2) * 64) & 511u;   // distinct column ranges for the warps of a quarter
    const long long t0 = clock64();
    if (mode == 0) {
      uint32_t v[16];
      for (int r = 0; r < rounds; ++r) {
        tmem_ld16(lane_addr + ((col0 + (uint32_t)(r & 3) * 16u) & 511u), v);
        tmem_ld_wait();
        sink += v[0] + v[15];
      }
    } else if (mode == 1) {
      uint32_t a[16], b[16], c[16], d[16];
      for (int r = 0; r < rounds; r += 4) {
        tmem_ld16(lane_addr + col0, a);
        tmem_ld16(lane_addr + col0 + 16, b);
        tmem_ld16(lane_addr + col0 + 32, c);
        tmem_ld16(lane_addr + col0 + 48, d);
        tmem_ld_wait();
        sink += a[0] + b[5] + c[7] + d[15];
      }
    } else if (mode == 2) {
      uint32_t v[32];
      for (int r = 0; r < rounds; r += 2) {
        tmem_ld32(lane_addr + col0 + (uint32_t)(r & 2) * 16u, v);
        tmem_ld_wait();
        sink += v[0] + v[31];
      }
    } else if (mode == 3) {
      uint32_t v[64];
      for (int r = 0; r < rounds; r += 4) {
        tmem_ld64(lane_addr + col0, v);
        tmem_ld_wait();
        sink += v[0] + v[63] + v[17];
      }
    } else if (mode == 4) {
      uint32_t v[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = lane + i;
      for (int r = 0; r < rounds; r += 2) {
        tmem_st32(lane_addr + col0 + (uint32_t)(r & 2) * 16u, v);
        tmem_st_wait();
      }
    } else if (mode == 5) {
      uint32_t v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = lane + i;
      for (int r = 0; r < rounds; r += 4) {
#pragma unroll
        for (int j = 0; j < 8; ++j) tmem_st8(lane_addr + col0 + 8u * j, v);
        tmem_st_wait();
      }
    }
    t_ld = clock64() - t0;
    __syncwarp();
    if (lane == 0) atomicAdd((int*)&ld_done, 1);
  } else if (warp == 16 && mma_mode != 0) {
    const int n = (mma_mode == 2 || mma_mode == 4) ? 256 : 128;
    const bool ts = mma_mode >= 3;
    const uint32_t idesc = idesc_16bit_f32(true, 128, n);
    const uint32_t a_lo = desc_lo_sw128(smem_u32(smem)), b_lo = desc_lo_sw128(smem_u32(smem) + 16384);
    const long long t0 = clock64();
    int issued = 0;
    // n_mma > 0: a fixed count; n_mma == 0: keep the pipe busy until the load warps are done
    while (true) {
      if (n_mma > 0 ? issued >= n_mma : ld_done >= n_ld_warps) break;
      // D at columns 256.. (N=128) or 256..511 (N=256): away from the columns the load warps read (0..255)
      if (ts)
        mma4_f16_ts_elect(256u, 200u, desc_from_lo(b_lo), idesc, 1u, 4);
      else
        mma4_f16_ss_elect(256u, desc_from_lo(a_lo), desc_from_lo(b_lo), idesc, 1u, 4);
      issued += 4;
    }
    t_issue = clock64() - t0;
    mma_commit_elect(&bar_done);
    mbar_wait(&bar_done, 0);
    t_mma = clock64() - t0;
    out[blockIdx.x].mma_issue = issued;   // overwritten below for n_mma > 0
  }
  __syncwarp();
  if (blockIdx.x == 0 || true) {
    if (threadIdx.x == 0) out[blockIdx.x].ld_cycles = t_ld;
    if (warp == 16 && lane == 0) {
      out[blockIdx.x].mma_cycles = t_mma;
      if (n_mma > 0) out[blockIdx.x].mma_issue = t_issue;
    }
    if (sink == 0x12345678u) out[blockIdx.x].sink = sink;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 16) tmem_dealloc<512>(0u);
}

int main() {
  Result* d;
  const int grid = 148;
  cudaMalloc(&d, grid * sizeof(Result));
  cudaFuncSetAttribute(ubench, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  std::vector<Result> h(grid);
  auto run = [&](const char* name, int mode, int nw, int rounds, int mma_mode, int n_mma) {
    cudaMemset(d, 0, grid * sizeof(Result));
    ubench<<<grid, 576, 64 * 1024>>>(mode, nw, rounds, mma_mode, n_mma, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("%s: CUDA error %s\n", name, cudaGetErrorString(e));
      exit(1);
    }
    cudaMemcpy(h.data(), d, grid * sizeof(Result), cudaMemcpyDeviceToHost);
    double ld = 0, mm = 0, mi = 0;
    for (auto& r : h) { ld += r.ld_cycles; mm += r.mma_cycles; mi += r.mma_issue; }
    ld /= grid; mm /= grid; mi /= grid;
    if (nw > 0) {
      // each round moves 16 columns x 32 lanes x 4 B = 2 KB per warp
      const double bytes = (double)rounds * 2048.0 * nw;
      printf("%-44s warps %2d mma_mode %d: %8.0f cyc for %d rounds -> %6.1f cyc/round, %6.1f B/cyc/SM", name, nw, mma_mode,
             ld, rounds, ld / rounds, bytes / ld);
      if (mma_mode) printf("   [mma issued meanwhile: %.0f, %.1f cyc/mma]", mi, mm / (mi > 0 ? mi : 1));
      printf("\n");
    } else {
      printf("%-44s mma_mode %d: %d MMAs issue %8.0f cyc (%.1f/mma), complete %8.0f cyc (%.1f/mma)\n", name, mma_mode, n_mma,
             mi, mi / n_mma, mm, mm / n_mma);
    }
  };
  run("warmup", 0, 4, 64, 0, 0);
  for (int nw : {1, 4, 8, 16}) {
    run("ld x16 + wait", 0, nw, 256, 0, 0);
    run("4 x ld x16, one wait", 1, nw, 256, 0, 0);
    run("ld x32 + wait", 2, nw, 256, 0, 0);
    run("ld x64 + wait", 3, nw, 256, 0, 0);
    run("st x32 + wait", 4, nw, 256, 0, 0);
    run("8 x st x8, one wait", 5, nw, 256, 0, 0);
  }
  run("mma SS N=128", 0, 0, 0, 1, 2048);
  run("mma SS N=256", 0, 0, 0, 2, 2048);
  run("mma TS N=128", 0, 0, 0, 3, 2048);
  run("mma TS N=256", 0, 0, 0, 4, 2048);
  for (int mm : {1, 3}) {
    run("ld x16 + wait, MMAs running", 0, 8, 256, mm, 0);
    run("4 x ld x16, one wait, MMAs running", 1, 8, 256, mm, 0);
    run("ld x64 + wait, MMAs running", 3, 8, 256, mm, 0);
    run("st x32 + wait, MMAs running", 4, 8, 256, mm, 0);
  }
  return 0;
}
