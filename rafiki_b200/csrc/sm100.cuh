// Inline-PTX building blocks for sm_100a: mbarrier, bulk async copies (TMA engine),
// tcgen05 MMA / TMEM, shared-memory matrix descriptors.
//
// Layout convention used by every kernel in this library for a tensor-core operand held in
// shared memory ("K-major, 128-byte swizzle"): a tile of R rows (R % 8 == 0) by 64 bf16
// K-elements is R consecutive 128-byte rows; the 16-byte chunk c (0..7) of row r is stored
// at chunk position (c ^ (r & 7)).  The tile base is 1024-byte aligned, so 8-row groups are
// 1024 bytes apart (the descriptor's stride-byte-offset).  Wider K is a sequence of such
// tiles ("atoms").  This is the layout a TMA tensor load with CU_TENSOR_MAP_SWIZZLE_128B
// produces, and the one pack_kmajor_sw128() in pack.cuh builds on the host for weights.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace sm100 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ------------------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// non-blocking probe
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// try_wait that lets the hardware park the warp for up to `ns` nanoseconds per attempt
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must trap (-> a CUDA error the host reports) instead of hanging
// the GPU.  The spin is cheap on issue slots (the warp sleeps inside try_wait) and gives up
// after ~2^22 attempts of up to 1 us each.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  uint32_t spins = 0;
#ifndef RFK_MBAR_HINT_NS
#define RFK_MBAR_HINT_NS 1000u
#endif
  while (!(RFK_MBAR_HINT_NS == 0u ? mbar_try_wait(bar, parity) : mbar_try_wait_hint(bar, parity, RFK_MBAR_HINT_NS))) {
    if (++spins > (1u << 22)) {
      printf("rafiki_b200: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", blockIdx.x,
             threadIdx.x, smem_u32(bar), parity);
      __trap();
    }
  }
}

// ------------------------------------------------------------------ proxies and fences
// generic-proxy smem writes -> visible to the async proxy (tcgen05.mma operand reads, bulk copies)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// ----------------------------------------------------------------- bulk async copy (TMA)
// 1-D global -> shared bulk copy, completion counted in bytes on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// 2-D tiled tensor load through a CUtensorMap (SASS: UTMALDG).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
          "r"(smem_u32(smem_dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// Warp-converged variants (see mma_f16_ss_elect): ALL 32 lanes execute the statement and
// elect.sync predicates the instruction itself on one lane.  Issuing these from inside an
// `if (lane == 0)` region makes the compiler wrap every instruction that takes uniform-register
// operands (UBLKCP, UTMALDG, the expect_tx arrive) in a serialising loop of a few hundred cycles.
__device__ __forceinline__ void bulk_g2s_expect_elect(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%3], %2;\n\t"
      "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n\t}" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_expect_elect(void* smem_dst, const void* tmap, int c0, int c1, uint32_t bytes,
                                                         uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%2], %5;\n\t"
      "@q cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}" ::
          "r"(smem_u32(smem_dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(bytes)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q mbarrier.arrive.shared::cta.b64 _, [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

// ------------------------------------------------------------------------------- TMEM
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {  // the same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns: thread i of the warp gets lane (base_lane + i).
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 8 consecutive 32-bit columns, registers -> TMEM: thread i writes lane (base_lane + i).
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]),
               "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
// 32 lanes x 16 consecutive 32-bit columns, registers -> TMEM.
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------ tcgen05.mma
// Shared-memory matrix descriptor, K-major operand, 128-byte swizzle (see header comment).
// bits [0,14) start address >> 4; [16,30) leading byte offset >> 4 (unused for swizzled
// K-major, 1); [32,46) stride byte offset >> 4 (1024 B between 8-row groups);
// [46,48) descriptor version 1 (Blackwell); [61,64) layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// The same descriptor split in two: the upper word is constant, the lower word carries the
// address, so advancing inside a swizzle atom is a 32-bit add.
__device__ __forceinline__ uint32_t desc_lo_sw128(uint32_t smem_addr) {
  return ((smem_addr & 0x3FFFFu) >> 4) | (1u << 16);
}
__device__ __forceinline__ uint64_t desc_from_lo(uint32_t lo) {
  constexpr uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | (uint64_t)lo;
}
// Instruction descriptor for kind::f16 with fp16 or bf16 A/B (both K-major), fp32 accumulate:
// [4,6) D format 1 = f32; [7,10) A format (0 = f16, 1 = bf16); [10,13) B format likewise;
// bit 15/16 A/B major 0 = K; [17,23) N >> 3; [24,29) M >> 4.
__host__ __device__ constexpr uint32_t idesc_16bit_f32(bool f16, int m, int n) {
  return (1u << 4) | ((f16 ? 0u : 1u) << 7) | ((f16 ? 0u : 1u) << 10) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]^T, issued by ONE thread for the CTA.
__device__ __forceinline__ void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Warp-converged variants: ALL 32 lanes execute the statement, elect.sync picks one lane and
// the instruction itself is predicated on it.  The surrounding code stays convergent, so the
// compiler keeps descriptors in uniform registers and emits no serialising loop per MMA.
__device__ __forceinline__ void mma_f16_ss_elect(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Up to four K steps (16 elements = 32 bytes = +2 on the descriptor each) of one swizzle atom in
// ONE statement: a single elect, one set of uniform operands, four predicated MMAs.  `nsteps`
// is 1..4; the first step accumulates iff `accumulate_first`, the others always do.
__device__ __forceinline__ void mma4_f16_ss_elect(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                   uint32_t accumulate_first, int nsteps) {
  asm volatile(
      "{\n\t.reg .pred p, q, q2, q3, q4, t;\n\t"
      ".reg .b64 a2, b2, a3, b3, a4, b4;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "setp.gt.and.s32 q2, %5, 1, q;\n\t"
      "setp.gt.and.s32 q3, %5, 2, q;\n\t"
      "setp.gt.and.s32 q4, %5, 3, q;\n\t"
      "add.u64 a2, %1, 2;\n\tadd.u64 b2, %2, 2;\n\t"
      "add.u64 a3, %1, 4;\n\tadd.u64 b3, %2, 4;\n\t"
      "add.u64 a4, %1, 6;\n\tadd.u64 b4, %2, 6;\n\t"
      "@q  tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "@q2 tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %3, t;\n\t"
      "@q3 tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %3, t;\n\t"
      "@q4 tcgen05.mma.cta_group::1.kind::f16 [%0], a4, b4, %3, t;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate_first), "r"(nsteps)
      : "memory");
}
// Same, with the A operand in tensor memory ("TS" form): A is 128 lanes x (K/2) 32-bit columns,
// two consecutive K elements per cell, so a 16-element K step is +8 columns; B stays in shared
// memory (+2 on the descriptor per step).
__device__ __forceinline__ void mma4_f16_ts_elect(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                                   uint32_t accumulate_first, int nsteps) {
  asm volatile(
      "{\n\t.reg .pred p, q, q2, q3, q4, t;\n\t"
      ".reg .b64 b2, b3, b4;\n\t"
      ".reg .b32 a2, a3, a4;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "setp.gt.and.s32 q2, %5, 1, q;\n\t"
      "setp.gt.and.s32 q3, %5, 2, q;\n\t"
      "setp.gt.and.s32 q4, %5, 3, q;\n\t"
      "add.u32 a2, %1, 8;\n\tadd.u64 b2, %2, 2;\n\t"
      "add.u32 a3, %1, 16;\n\tadd.u64 b3, %2, 4;\n\t"
      "add.u32 a4, %1, 24;\n\tadd.u64 b4, %2, 6;\n\t"
      "@q  tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "@q2 tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %3, t;\n\t"
      "@q3 tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %3, t;\n\t"
      "@q4 tcgen05.mma.cta_group::1.kind::f16 [%0], [a4], b4, %3, t;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate_first), "r"(nsteps)
      : "memory");
}
__device__ __forceinline__ void mma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
// Arrive on an mbarrier once every previously issued tcgen05.mma of this thread has
// completed (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}


// ------------------------------------------------------------------ CTA pairs (cluster of two, cta_group::2)
// Two CTAs on the SMs of one TPC act as one 256-row tile: ONE thread of the leader CTA (cluster
// rank 0) issues each tcgen05.mma; every CTA supplies its own 128 rows of A (tensor memory) and of
// D (tensor memory) and HALF of the B operand's N rows from its own shared memory, at the same
// offsets in both CTAs.  Completion is multicast to the same mbarrier offset in both CTAs.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {   // every thread of both CTAs
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address -> the same offset in CTA `rank` of the cluster (shared::cluster window)
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// Arrive on an mbarrier of any CTA of the cluster (address from map_to_cta).  Default semantics
// (release at CTA scope): what crosses the pair here is tensor-memory state, ordered by the tcgen05
// fences around the arrive and the wait.  The .release.cluster / .acquire.cluster forms compile to
// MEMBAR.ALL.GPU and CCTL.IVALL (an L1 invalidate) per call and made every hand-off several times slower.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_slot) {  // one full warp, the same warp index in both CTAs
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// mma4_f16_ts_elect for a CTA pair (M = 256): leader CTA only
__device__ __forceinline__ void mma4_f16_ts_pair_elect(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                                        uint32_t accumulate_first, int nsteps) {
  asm volatile(
      "{\n\t.reg .pred p, q, q2, q3, q4, t;\n\t"
      ".reg .b64 b2, b3, b4;\n\t"
      ".reg .b32 a2, a3, a4;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "setp.gt.and.s32 q2, %5, 1, q;\n\t"
      "setp.gt.and.s32 q3, %5, 2, q;\n\t"
      "setp.gt.and.s32 q4, %5, 3, q;\n\t"
      "add.u32 a2, %1, 8;\n\tadd.u64 b2, %2, 2;\n\t"
      "add.u32 a3, %1, 16;\n\tadd.u64 b3, %2, 4;\n\t"
      "add.u32 a4, %1, 24;\n\tadd.u64 b4, %2, 6;\n\t"
      "@q  tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "@q2 tcgen05.mma.cta_group::2.kind::f16 [%0], [a2], b2, %3, t;\n\t"
      "@q3 tcgen05.mma.cta_group::2.kind::f16 [%0], [a3], b3, %3, t;\n\t"
      "@q4 tcgen05.mma.cta_group::2.kind::f16 [%0], [a4], b4, %3, t;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate_first), "r"(nsteps)
      : "memory");
}
// completion of the leader's earlier MMAs -> the mbarrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void mma_commit_pair_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t.reg .b16 m;\n\t"
      "mov.b16 m, 3;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(
          smem_u32(bar))
      : "memory");
}

// Four K steps of 16, always (the hot path of the layer-1 issuer: its loop is a serial chain on the critical
// path).  Executed by the whole warp so that every operand stays in uniform registers; one elected lane
// issues.  The descriptors of steps 2..4 differ in their low word only (start address + 32 bytes per step).
__device__ __forceinline__ void mma4_f16_ts_pair_whole(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_desc_lo, uint32_t idesc,
                                                        uint32_t accumulate_first) {
  constexpr uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);      // as desc_from_lo
  asm volatile(
      "{\n\t.reg .pred p, q, t;\n\t"
      ".reg .b64 b1, b2, b3, b4;\n\t"
      ".reg .b32 l2, l3, l4, a2, a3, a4;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "add.u32 l2, %2, 2;\n\tadd.u32 l3, %2, 4;\n\tadd.u32 l4, %2, 6;\n\t"
      "add.u32 a2, %1, 8;\n\tadd.u32 a3, %1, 16;\n\tadd.u32 a4, %1, 24;\n\t"
      "mov.b64 b1, {%2, %5};\n\tmov.b64 b2, {l2, %5};\n\tmov.b64 b3, {l3, %5};\n\tmov.b64 b4, {l4, %5};\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], b1, %3, p;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], [a2], b2, %3, t;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], [a3], b3, %3, t;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], [a4], b4, %3, t;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "r"(b_desc_lo), "r"(idesc), "r"(accumulate_first), "r"(hi)
      : "memory");
}

// Programmatic dependent launch (a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization):
// griddep_launch_dependents lets the NEXT kernel of the stream start its prologue as soon as every CTA of this
// grid has executed it (or exited); griddep_wait blocks until the PREVIOUS grid has completed and its memory
// is visible.  Both are no-ops for an ordinary launch.
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

}  // namespace sm100
