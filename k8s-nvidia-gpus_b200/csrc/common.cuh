// Shared device-side primitives for the b200sd kernels (sm_100a only).
//
// Thin inline-PTX wrappers for mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (UMMA + TMEM)
// plus descriptor builders.  No CUTLASS dependency: bit layouts follow the PTX ISA tables for
// the tcgen05 shared-memory matrix descriptor and the kind::f16 instruction descriptor.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace b200sd {

// 16-bit activation / weight element type of every kernel.  Default: IEEE fp16 — the reference's own CUDA dtype
// (TORCH_DTYPE = float16, configmap.yaml:31): its 11-bit significand keeps the CFG-amplified per-step latent error
// inside the 1e-2 parity tolerance, at the same tcgen05 kind::f16 throughput as bf16.  -DB200SD_ACT_BF16 builds the
// bfloat16 variant (8-bit significand, fp32 exponent range).
#ifdef B200SD_ACT_BF16
typedef __nv_bfloat16 act_t;
static constexpr uint32_t kUmmaFmt = 1;  // kind::f16 A/B format field: 0 = f16, 1 = bf16
static constexpr int kActDtypeCode = 0;
#else
typedef __half act_t;
static constexpr uint32_t kUmmaFmt = 0;
static constexpr int kActDtypeCode = 1;
#endif

// Every mbarrier wait is bounded: a protocol bug must trap (error surfaces to the host), not hang the GPU.
#ifndef B200SD_WAIT_TIMEOUT_CYCLES
#define B200SD_WAIT_TIMEOUT_CYCLES 4000000000LL
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Cold path of every bounded wait, kept out of line so that wait sites stay a handful of instructions.
static __device__ __noinline__ void mbar_timeout_trap(uint32_t bar, uint32_t parity) {
  printf("b200sd: mbarrier wait timeout (block %d,%d,%d thread %d bar 0x%x parity %u)\n", blockIdx.x, blockIdx.y,
         blockIdx.z, threadIdx.x, bar, parity);
  __trap();
}
// One poll with a suspend-time hint: the hardware may park the thread while the phase is incomplete, so a waiting
// warp burns almost no issue slots of the SM sub-partition it shares with working warps (long waits).
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(done)
      : "r"(bar), "r"(parity), "r"(0x989680u)
      : "memory");
  return done != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t polls = 0;
  long long t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    // the wall-clock guard is evaluated once per 1024 polls
    if (((++polls) & 1023u) == 0u) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > B200SD_WAIT_TIMEOUT_CYCLES) mbar_timeout_trap(bar, parity);
    }
  }
}
// Latency-critical variant: plain polling (no suspend hint -> no NANOSLEEP wake-up latency), same bounded guard.
__device__ __forceinline__ void mbar_wait_spin(uint32_t bar, uint32_t parity) {
  uint32_t done;
  uint32_t polls = 0;
  long long t0 = 0;
  for (;;) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.b32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return;
    if (((++polls) & 4095u) == 0u) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > B200SD_WAIT_TIMEOUT_CYCLES) mbar_timeout_trap(bar, parity);
    }
  }
}

// ------------------------------------------------------------------ programmatic dependent launch
// A kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start (on SMs the previous kernel has
// left) before that kernel has finished: its set-up - barrier init, TMEM allocation, descriptor prefetch - then runs
// under the previous kernel's tail.  pdl_wait() returns once every earlier kernel of the stream has completed and its
// writes are visible: NOTHING produced inside the step may be read, and no global memory written, before it.
// pdl_launch() lets the NEXT kernel's blocks be scheduled as soon as this grid's blocks have all issued it or exited.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// generic-proxy writes to shared memory -> visible to the async proxy (TMA / UMMA operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ------------------------------------------------------------------ TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// smem -> global tile store (bulk async group); OOB parts of the box are clipped by the TMA unit
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tm)),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed bulk stores have finished READING shared memory (the staging tile may be overwritten)
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... all but the most recently committed one
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ CTA pairs (cta_group::2)
// In a 2-CTA cluster the even CTA leads: it owns the "full" barriers and issues the MMAs.  Clearing bit 24 of a
// shared::cluster address turns the peer's view of a barrier into the leader's copy of it.
static constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t rank) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(bar),
      "r"(rank)
      : "memory");
}
// TMA loads issued by either CTA of a pair; the transaction bytes are credited to the LEADER's barrier
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_pair(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1,
                                                 int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4, %5, %6}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem of both CTAs] (+)= A (128 rows from each CTA) * B (N/2 rows from each CTA); issued by the leader only
__device__ __forceinline__ void umma_f16_ss_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs once all previously issued MMAs have completed
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"(static_cast<uint16_t>(3))
      : "memory");
}

// ------------------------------------------------------------------ warp-converged issue ("_w" variants)
// tcgen05.mma / tcgen05.commit / TMA take their operands from uniform registers.  When a single lane runs the issue
// loop inside a divergent branch the compiler cannot keep anything uniform and wraps EVERY such instruction in an
// elect / R2UR "waterfall" (20-30 SASS instructions per MMA - more than the tensor pipe needs to execute a small
// tile).  The issue loops therefore run with the WHOLE warp converged: addresses, descriptors and ring counters stay
// on the uniform datapath and only the instruction itself is predicated on the elected lane.
__device__ __forceinline__ uint32_t warp_idx_uniform() { return __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0); }
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t is_leader;
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "elect.sync _|e, 0xffffffff;\n"
      "selp.b32 %0, 1, 0, e;\n"
      "}\n"
      : "=r"(is_leader));
  return is_leader;
}
__device__ __forceinline__ void umma_f16_ss_w(uint32_t leader, uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi,
                                               uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p, e;\n"
      ".reg .b64 da, db;\n"
      "mov.b64 da, {%1, %2};\n"
      "mov.b64 db, {%3, %4};\n"
      "setp.ne.b32 p, %6, 0;\n"
      "setp.ne.b32 e, %7, 0;\n"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate), "r"(leader)
      : "memory");
}
// A operand in TMEM (lane = row, each 32-bit column holds two consecutive K elements), B from shared memory
__device__ __forceinline__ void umma_f16_ts_w(uint32_t leader, uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo,
                                               uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p, e;\n"
      ".reg .b64 db;\n"
      "mov.b64 db, {%2, %3};\n"
      "setp.ne.b32 p, %5, 0;\n"
      "setp.ne.b32 e, %6, 0;\n"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void umma_commit_w(uint32_t leader, uint32_t bar) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %1, 0;\n"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
      "}\n" ::"r"(bar),
      "r"(leader)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx_w(uint32_t leader, uint32_t bar, uint32_t bytes) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %2, 0;\n"
      "@e mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
      "}\n" ::"r"(bar),
      "r"(bytes), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_w(uint32_t leader, uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0,
                                               int c1, int c2, int c3) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %7, 0;\n"
      "@e cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];\n"
      "}\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(leader)
      : "memory");
}

__device__ __forceinline__ void tma_load_2d_w(uint32_t leader, uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0,
                                               int c1) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %5, 0;\n"
      "@e cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
      "}\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1), "r"(leader)
      : "memory");
}
// L2 prefetch of one box of a 2-D tensor map (no shared-memory destination, no barrier)
__device__ __forceinline__ void tma_prefetch_l2_2d_w(uint32_t leader, const CUtensorMap* tm, int c0, int c1) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %3, 0;\n"
      "@e cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];\n"
      "}\n" ::"l"(reinterpret_cast<uint64_t>(tm)),
      "r"(c0), "r"(c1), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair_w(uint32_t leader, uint32_t dst, const CUtensorMap* tm, uint32_t bar,
                                                    int c0, int c1) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %5, 0;\n"
      "@e cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4}], [%2];\n"
      "}\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_pair_w(uint32_t leader, uint32_t dst, const CUtensorMap* tm, uint32_t bar,
                                                    int c0, int c1, int c2, int c3) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %7, 0;\n"
      "@e cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4, %5, %6}], [%2];\n"
      "}\n" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void umma_f16_ss_pair_w(uint32_t leader, uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi,
                                                    uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p, e;\n"
      ".reg .b64 da, db;\n"
      "mov.b64 da, {%1, %2};\n"
      "mov.b64 db, {%3, %4};\n"
      "setp.ne.b32 p, %6, 0;\n"
      "setp.ne.b32 e, %7, 0;\n"
      "@e tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair_w(uint32_t leader, uint32_t bar) {
  asm volatile(
      "{\n"
      ".reg .pred e;\n"
      "setp.ne.b32 e, %1, 0;\n"
      "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %2;\n"
      "}\n" ::"r"(bar),
      "r"(leader), "h"(static_cast<uint16_t>(3))
      : "memory");
}

// ------------------------------------------------------------------ tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], act_t inputs, fp32 accumulate; issued by ONE thread.
__device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// 32 lanes x 64 consecutive fp32 columns -> 64 registers per thread.
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&v)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]), "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = TMEM lane).
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
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
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ descriptors
// Shared-memory matrix descriptor (64 bit):
//   [0,14) start address >> 4 | [16,30) leading byte offset >> 4 | [32,46) stride byte offset >> 4
//   [46,48) version = 1 (sm_100) | [49,52) base offset | [61,64) swizzle: 0 none, 2 128B, 4 64B, 6 32B
// K-major, SWIZZLE_128B, 64 act_t (128 B) per row: rows at 128 B pitch, 8-row groups 1024 B apart (SBO),
// LBO unused.  MN-major, SWIZZLE_128B: 64 MN-elements (128 B) contiguous, 8 K-rows at 128 B pitch,
// next 8 K-rows SBO apart, next 64 MN-elements LBO apart.
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

// Instruction descriptor for kind::f16, act_t x act_t -> fp32:
//   [4,6) D format (1 = f32) | [7,10) A format (1 = act_t) | [10,13) B format | 15 A major | 16 B major (1 = MN-major)
//   [17,23) N >> 3 | [24,29) M >> 4
__device__ __forceinline__ uint32_t make_idesc_f16(uint32_t M, uint32_t N, uint32_t a_mn_major, uint32_t b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4;
  d |= kUmmaFmt << 7;
  d |= kUmmaFmt << 10;
  d |= (a_mn_major & 1u) << 15;
  d |= (b_mn_major & 1u) << 16;
  d |= ((N >> 3) & 0x3Fu) << 17;
  d |= ((M >> 4) & 0x1Fu) << 24;
  return d;
}

// ------------------------------------------------------------------ small math / packing helpers
#ifdef B200SD_ACT_BF16
__device__ __forceinline__ uint32_t pack_act2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_act2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
__device__ __forceinline__ act_t to_act(float x) { return __float2bfloat16_rn(x); }
__device__ __forceinline__ float from_act(act_t x) { return __bfloat162float(x); }
#else
__device__ __forceinline__ uint32_t pack_act2(float lo, float hi) {
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_act2(uint32_t u) {
  __half2 v = *reinterpret_cast<__half2*>(&u);
  return __half22float2(v);
}
__device__ __forceinline__ act_t to_act(float x) { return __float2half_rn(x); }
__device__ __forceinline__ float from_act(act_t x) { return __half2float(x); }
#endif
// Activations for the epilogues: two MUFU ops each, no IEEE division, no branches (these run per output element of
// the short-K GEMMs and of every GroupNorm, where the epilogue - not the tensor pipe - sets the pace).
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// x * sigmoid(x); x -> -inf: ex2 -> +inf, rcp -> 0, result -0
__device__ __forceinline__ float silu_f(float x) { return x * rcp_approx(1.f + ex2_approx(-1.4426950408889634f * x)); }
// x * sigmoid(x) with ONE MUFU op: sigmoid(x) = 0.5 + 0.5 tanh(x / 2) (tanh.approx: relative error 2^-11, i.e. an absolute
// error below 2.5e-4 |x| - under the 16-bit rounding of an O(1) activation)
__device__ __forceinline__ float silu_tanh_f(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  return x * fmaf(0.5f, t, 0.5f);
}
__device__ __forceinline__ float quick_gelu_f(float x) {
  return x * rcp_approx(1.f + ex2_approx(-1.702f * 1.4426950408889634f * x));
}
// exact (erf) GELU = x * Phi(x) with ONE MUFU op per element: log2 Phi(-a), a = |x|, is a smooth function that a
// degree-6 polynomial through (0, -1) fits on [0, 6] to 2.4e-7 of a * erfc (weighted minimax; beyond 6, Phi(-a) < 1e-9),
// so  h = Phi(-a) = 2^q(a)  costs six FMAs and an ex2, and  x Phi(x) = 0.5 x + a (0.5 - h)  for either sign.  |error| of
// the result < 5e-7 over [-10, 10] evaluated in fp32 - the same as the Abramowitz & Stegun 7.1.26 form it replaces
// (which needed a reciprocal and an exponential per element: the GEGLU epilogues were bound by the MUFU pipe).
static constexpr float kGeluQ1 = -1.1511470079421997f, kGeluQ2 = -0.458916038274765f, kGeluQ3 = -0.05323765426874161f,
                       kGeluQ4 = 0.007977096363902092f, kGeluQ5 = -0.0007397594163194299f,
                       kGeluQ6 = 2.9910861485404894e-05f;
__device__ __forceinline__ float gelu_erf_f(float x) {
  const float a = fabsf(x), c = fminf(a, 6.f);
  float q = fmaf(kGeluQ6, c, kGeluQ5);
  q = fmaf(q, c, kGeluQ4);
  q = fmaf(q, c, kGeluQ3);
  q = fmaf(q, c, kGeluQ2);
  q = fmaf(q, c, kGeluQ1);
  q = fmaf(q, c, -1.f);
  const float h = ex2_approx(q);  // Phi(-|x|)
  return fmaf(a, 0.5f - h, 0.5f * x);
}

// ------------------------------------------------------------------ order-independent statistics accumulation
// GroupNorm sums are gathered from many CTAs / warps.  fp32 atomics make the result depend on arrival order (images
// differed by a few LSB between identical runs), so the partial sums are accumulated as 64-bit INTEGERS instead:
// integer addition is associative, the total is the same whichever producer arrives first.  A fp32 partial p becomes
// fixed point with LSB 2^-24, split so that no carry ever crosses words:
//     p * 2^24 = H * 2^31 + L,  0 <= L < 2^31      acc[0] += H,  acc[1] += L     (two independent 64-bit sums)
// Both conversions are exact up to the 2^-24 LSB (power-of-two scalings, floor, and a remainder that only drops high
// bits).  Range: |total| < 2^63 * 128 for H and 2^32 partials for L - far beyond any tensor here.
// One (image, group) statistics slot = 4 such words {sum H, sum L, sum-of-squares H, sum-of-squares L} = 32 bytes
// (B200SD_GN_STATS_WORDS 4-byte words in b200sd.h).
static constexpr int kGnSlotWords64 = 4;
__device__ __forceinline__ void fixed_split(float p, long long& H, long long& L) {
  const float h = floorf(p * 0.0078125f);  // p / 128, exact
  H = __float2ll_rn(h);
  const float r = fmaf(-128.f, h, p);      // 0 <= r < 128, exact (only high bits are removed)
  L = __float2ll_rn(r * 16777216.f);
}
__device__ __forceinline__ void fixed_add_global(unsigned long long* slot2, float p) {
  long long H, L;
  fixed_split(p, H, L);
  atomicAdd(slot2, static_cast<unsigned long long>(H));  // result unused: compiles to fire-and-forget RED.64
  atomicAdd(slot2 + 1, static_cast<unsigned long long>(L));
}
__device__ __forceinline__ double fixed_value(const unsigned long long* slot2) {
  return static_cast<double>(static_cast<long long>(slot2[0])) * 128.0 +
         static_cast<double>(static_cast<long long>(slot2[1])) * 5.9604644775390625e-08;  // 2^-24
}

// ------------------------------------------------------------------ packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2)
// One instruction for two lanes' worth of fp32 work: the epilogues are issue/latency-bound, not FMA-pipe-bound, so
// halving the instruction count of their element-wise chains is what counts.
__device__ __forceinline__ void ffma2(float a0, float a1, float b0, float b1, float c0, float c1, float& x0, float& x1) {
  asm("{\n.reg .b64 va, vb, vc, vd;\nmov.b64 va, {%2, %3};\nmov.b64 vb, {%4, %5};\nmov.b64 vc, {%6, %7};\n"
      "fma.rn.f32x2 vd, va, vb, vc;\nmov.b64 {%0, %1}, vd;\n}\n"
      : "=f"(x0), "=f"(x1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ void fmul2(float a0, float a1, float b0, float b1, float& x0, float& x1) {
  asm("{\n.reg .b64 va, vb, vd;\nmov.b64 va, {%2, %3};\nmov.b64 vb, {%4, %5};\nmul.rn.f32x2 vd, va, vb;\n"
      "mov.b64 {%0, %1}, vd;\n}\n"
      : "=f"(x0), "=f"(x1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ void fadd2(float a0, float a1, float b0, float b1, float& x0, float& x1) {
  asm("{\n.reg .b64 va, vb, vd;\nmov.b64 va, {%2, %3};\nmov.b64 vb, {%4, %5};\nadd.rn.f32x2 vd, va, vb;\n"
      "mov.b64 {%0, %1}, vd;\n}\n"
      : "=f"(x0), "=f"(x1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
// (v0, v1) *= gelu_erf(g0), gelu_erf(g1): the pair form of gelu_erf_f (same formula, same constants)
__device__ __forceinline__ void mul_gelu_erf2(float& v0, float& v1, float g0, float g1) {
  const float a0 = fabsf(g0), a1 = fabsf(g1);
  const float c0 = fminf(a0, 6.f), c1 = fminf(a1, 6.f);
  float q0, q1, s0, s1, w0, w1;
  ffma2(c0, c1, kGeluQ6, kGeluQ6, kGeluQ5, kGeluQ5, q0, q1);
  ffma2(q0, q1, c0, c1, kGeluQ4, kGeluQ4, q0, q1);
  ffma2(q0, q1, c0, c1, kGeluQ3, kGeluQ3, q0, q1);
  ffma2(q0, q1, c0, c1, kGeluQ2, kGeluQ2, q0, q1);
  ffma2(q0, q1, c0, c1, kGeluQ1, kGeluQ1, q0, q1);
  ffma2(q0, q1, c0, c1, -1.f, -1.f, q0, q1);
  const float h0 = ex2_approx(q0), h1 = ex2_approx(q1);  // Phi(-|g|)
  ffma2(h0, h1, -1.f, -1.f, 0.5f, 0.5f, s0, s1);         // 0.5 - h
  fmul2(g0, g1, 0.5f, 0.5f, w0, w1);
  ffma2(a0, a1, s0, s1, w0, w1, w0, w1);                 // g Phi(g) = 0.5 g + |g| (0.5 - h)
  fmul2(v0, v1, w0, w1, v0, v1);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace b200sd
