// Flash attention forward on tcgen05 (sm_100a): out = softmax(Q K^T * scale [+ causal]) V per (batch, head).
// SURVEY §2.1 K5 (self-attention), K6 (cross-attention, nk = 77), K15 (CLIP causal attention).
//
// One CTA per (up to two 128-query tiles, head, batch); keys are walked in tiles of 64; 20 warps:
//   warp 0            : TMA producer.  Q/K/V are read straight out of the projection GEMM outputs through 4-D
//                       tensor maps {d, head, token, batch}; the box is 64 wide in d, so head dims that are not a
//                       multiple of 64 (40, 80, 160) and ragged token counts are zero-filled by the TMA unit —
//                       no padded copies of Q/K/V exist in HBM.  K/V tiles are shared by both query tiles.
//   warps 1, 3        : tcgen05.mma issuer of query tile 0 / 1 (whole warp converged, one elected lane issues).
//                       S_t(j) = Q_t K(j)^T (K-major A and B) goes to TMEM buffer j & 1 of query tile t;
//                       O_t,w += P_t(j) V(j) with P_t(j) (16-bit, shared memory) as the K-major A operand and V
//                       consumed in place as an MN-major B operand, into the accumulator of parity w = j & 1.
//                       S_t(j+2) is issued as soon as its owner has copied S_t(j) to registers (s_free), i.e. while
//                       the exponentials of tile j are still being computed.
//   warp 2            : TMEM allocator.
//   warps 4..19       : four softmax warpgroups, two per query tile (thread = query row = TMEM lane).  The two
//                       warpgroups of a query tile take ALTERNATE KV tiles (even / odd), each with its own S buffer,
//                       P buffer, lazily updated reference and O accumulator - no per-tile exchange between them -
//                       and four resident warps per SM sub-partition keep the MUFU pipe fed.  Per tile: tcgen05.ld of
//                       the 64 scores -> release S -> row max -> p = exp2(s * scale - ref) (reference moved, with a
//                       rescale of O in TMEM, only when the tile max overshoots it by 2^8) -> 16-bit P -> swizzled
//                       shared memory.  At the end the two accumulators of a query tile are merged like split-KV
//                       partials: O = (f_e O_e + f_o O_o) / (f_e l_e + f_o l_o), f_x = 2^(ref_x - max ref).
//   TMEM (512 columns): d <= 64: 2 query tiles x (2 S buffers x 64 + 2 accumulators x 64); larger head dims: 1 query
//   tile x (2 x 64 + 2 x 192).
//   Row sums are kept in registers (64 FADDs per tile and thread).  Both ways of getting them from the tensor pipe
//   were measured slower on the 4096-token d = 40 call: a ones column written into every V tile (0.794 ms - the
//   softmax warps then wait for V to land) and an extra P x ONES product with N = 16 (0.843 ms) against 0.735 ms -
//   every tcgen05.mma re-reads its 128 x 16 A slice (4 KB) from shared memory whatever N is, and the shared-memory
//   port (operand reads + P stores + TMA writes) is as loaded as the MUFU pipe in this kernel.
#include <stdlib.h>

#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kAttThreads = 640;  // 4 control warps + 4 softmax warpgroups
static constexpr int kBQ = 128;  // query rows per tile (UMMA M)
static constexpr int kBKV = 64;  // keys per tile
static constexpr uint32_t kQAtom = 128 * 128;   // Q / P: 128 rows x 64 elements (one swizzle-128B atom column)
static constexpr uint32_t kKVAtom = kBKV * 128;  // K / V: 64 rows x 64 elements
static constexpr int kMaxStages = 4;
// mbarrier indices inside the kernel's barrier block ([t] = query tile, [w] = KV-tile parity / warpgroup)
enum : int {
  kBarQFull = 0,
  kBarQFree = 1,                      // the item's S products have completed: Q may be overwritten
  kBarOFree = 2,                      // [t]     the merge has read both accumulators of query tile t
  kBarSFull = 4,                      // [t][w]  S tile written by the tensor pipe
  kBarSFree = kBarSFull + 4,          // [t][w]  S tile copied to registers
  kBarPFull = kBarSFree + 4,          // [t][w]  P tile in shared memory
  kBarPVDone = kBarPFull + 4,         // [t][w]  P V accumulated
  kBarKFull = kBarPVDone + 4,         // [stage]
  kBarVFull = kBarKFull + kMaxStages,
  kBarKEmpty = kBarVFull + kMaxStages,
  kBarVEmpty = kBarKEmpty + kMaxStages,
  kNumBars = kBarVEmpty + kMaxStages
};

struct alignas(64) AttnDev {
  CUtensorMap tmQ, tmK, tmV;
  act_t* out;
  int batch, heads, nq, nk, d;
  int ldo;
  int nd;         // 64-wide d atoms
  int dN;         // UMMA N of the PV product / K extent of QK^T (d rounded up to 16)
  int stages;     // K/V ring depth
  int qtiles;     // query tiles per CTA (1 or 2)
  int o_col;      // first TMEM column of the O accumulators
  int o_stride;   // TMEM column stride between accumulators (two per query tile: even / odd KV tiles)
  int p_col;      // >= 0: P travels through TMEM (A operand of P V read from tensor memory): first column of the
                  // per-query-tile P buffers (32 columns each, 16-bit pairs); -1: P goes through shared memory
  int causal;
  int nqblk;      // query blocks (of qtiles * 128 rows) per (batch, head)
  int items;      // work items = batch * heads * nqblk
  float scale_log2;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// (x0, x1) = (a0, a1) * s + c as one fma.rn.f32x2
__device__ __forceinline__ void fma2(float a0, float a1, float s, float c, float& x0, float& x1) {
  asm("{\n"
      ".reg .b64 va, vs, vc, vd;\n"
      "mov.b64 va, {%2, %3};\n"
      "mov.b64 vs, {%4, %4};\n"
      "mov.b64 vc, {%5, %5};\n"
      "fma.rn.f32x2 vd, va, vs, vc;\n"
      "mov.b64 {%0, %1}, vd;\n"
      "}\n"
      : "=f"(x0), "=f"(x1)
      : "f"(a0), "f"(a1), "f"(s), "f"(c));
}

__global__ void __launch_bounds__(kAttThreads, 1) attention_kernel(const __grid_constant__ AttnDev p) {
  extern __shared__ uint8_t smem_raw[];
  // all mbarriers in one block: one base address, every barrier a constant byte offset from it
  __shared__ __align__(8) uint64_t bars[kNumBars];
  __shared__ uint32_t tmem_base_s;
  const uint32_t bar0 = smem_u32(bars);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_bytes = p.nd * kQAtom;
  const uint32_t kv_bytes = p.nd * kKVAtom;
  const uint32_t sQ = base;                          // [qtiles][nd atoms of 16 KB]
  const uint32_t sK = sQ + p.qtiles * q_bytes;       // [stages][nd atoms of 8 KB]
  const uint32_t sV = sK + p.stages * kv_bytes;      // [stages][nd atoms of 8 KB]
  const uint32_t sP = sV + p.stages * kv_bytes;      // [qtiles][2 buffers][16 KB]  (absent when P lives in TMEM)
  const uint32_t sX = sP + (p.p_col >= 0 ? 0 : p.qtiles * 2 * kQAtom);  // [qtiles][2 warpgroups][ref, sum][128] floats

  // PERSISTENT: the CTA walks work items (a block of 1-2 query tiles of one (batch, head)) with stride gridDim.x.
  // TMEM, barriers and the K/V ring are set up once; every barrier parity below comes from a running count of that
  // barrier's completions, so the Q load and the first Q K^T of the next item run under the last tiles, the merge and
  // the store of the current one, and there is no per-item allocation, initialisation or drain.
  // item -> (query block, head, batch), query block fastest (concurrent CTAs share a (batch, head)'s K/V in L2)
  auto item_geom = [&](int item, int& q0, int& head, int& b, int& ntiles, int& T) {
    const int qb = item % p.nqblk;
    const int bh = item / p.nqblk;
    head = bh % p.heads;
    b = bh / p.heads;
    q0 = qb * kBQ * p.qtiles;
    ntiles = (p.qtiles == 2 && q0 + kBQ < p.nq) ? 2 : 1;  // query tiles this item really has
    int kv_end = p.nk;
    if (p.causal) kv_end = min(p.nk, q0 + kBQ * ntiles);  // keys beyond the last query row are fully masked
    T = (kv_end + kBKV - 1) / kBKV;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQ);
    tma_prefetch_desc(&p.tmK);
    tma_prefetch_desc(&p.tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init((bar0 + 8 * kBarQFull), 1);
    mbar_init((bar0 + 8 * kBarQFree), p.qtiles);  // one commit per MMA issuer: the item's S products have read Q
    for (int t = 0; t < 2; ++t) {
      mbar_init((bar0 + 8 * (kBarOFree + t)), 256);  // both softmax warpgroups of the tile have read O out
      for (int i = 0; i < 2; ++i) {
        mbar_init((bar0 + 8 * (kBarSFull + 2 * (t) + (i))), 1);
        mbar_init((bar0 + 8 * (kBarPFull + 2 * (t) + (i))), 128);
        mbar_init((bar0 + 8 * (kBarSFree + 2 * (t) + (i))), 128);
        mbar_init((bar0 + 8 * (kBarPVDone + 2 * (t) + (i))), 1);
      }
    }
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init((bar0 + 8 * (kBarKFull + (s))), 1);
      mbar_init((bar0 + 8 * (kBarVFull + (s))), 1);
      mbar_init((bar0 + 8 * (kBarKEmpty + (s))), p.qtiles);  // one tcgen05.commit per MMA issuer warp (an issuer whose
      mbar_init((bar0 + 8 * (kBarVEmpty + (s))), p.qtiles);  // query tile does not exist in an item still arrives)
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(smem_u32(&tmem_base_s), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_s, 0);  // warp-uniform for the issue loops
  pdl_launch();  // set-up done (common.cuh: programmatic dependent launch)
  // TMEM columns: S_t[buf] at (2t + buf) * 64; accumulator of (query tile t, tile parity w) at o_col + (2t + w) * o_stride

  if (warp < 4) {
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (whole warp converged, one
    // elected lane issues - see the "_w" note in common.cuh)
    pdl_wait();  // q / k / v come from the previous kernels of the step
    const uint32_t leader = elect_one();
    int st = 0;
    uint32_t ph = 1;
    int n = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n) {
      int q0, head, b, ntiles, T;
      item_geom(item, q0, head, b, ntiles, T);
      if (n > 0) mbar_wait((bar0 + 8 * kBarQFree), (n - 1) & 1);  // the previous item's S products are done with Q
      mbar_arrive_expect_tx_w(leader, (bar0 + 8 * kBarQFull), ntiles * q_bytes);
      for (int t = 0; t < ntiles; ++t)
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d_w(leader, sQ + t * q_bytes + a * kQAtom, &p.tmQ, (bar0 + 8 * kBarQFull), a * 64, head, q0 + t * kBQ, b);
      for (int j = 0; j < T; ++j) {
        mbar_wait((bar0 + 8 * (kBarKEmpty + (st))), ph);
        mbar_arrive_expect_tx_w(leader, (bar0 + 8 * (kBarKFull + (st))), kv_bytes);
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d_w(leader, sK + st * kv_bytes + a * kKVAtom, &p.tmK, (bar0 + 8 * (kBarKFull + (st))), a * 64, head, j * kBKV, b);
        mbar_wait((bar0 + 8 * (kBarVEmpty + (st))), ph);
        mbar_arrive_expect_tx_w(leader, (bar0 + 8 * (kBarVFull + (st))), kv_bytes);
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d_w(leader, sV + st * kv_bytes + a * kKVAtom, &p.tmV, (bar0 + 8 * (kBarVFull + (st))), a * 64, head, j * kBKV, b);
        if (++st == p.stages) {
          st = 0;
          ph ^= 1u;
        }
      }
    }
  } else if ((warp == 1 || warp == 3) && (warp >> 1) < p.qtiles) {
    const int t = warp >> 1;  // warp 1 issues for query tile 0, warp 3 for query tile 1
    // ---------------------------------------------------------------- MMA issuer of query tile t
    // Whole warp converged; descriptors are built once and only their low words (16-byte units) move.  K and V
    // stages are consumed strictly in order, so each ring is tracked by a running (stage, phase) pair.
    const uint32_t leader = elect_one();
    const uint32_t idesc_s = make_idesc_f16(kBQ, kBKV, 0, 0);  // S[128 x 64] = Q (K-major) x K (K-major)
    const uint32_t idesc_o = make_idesc_f16(kBQ, p.dN, 0, 1);  // O[128 x dN] += P (K-major) x V (MN-major)
    const int ksteps_s = p.dN / 16;
    const uint64_t dQ = make_smem_desc_sw128(sQ + t * q_bytes, 16, 1024);
    const uint64_t dK = make_smem_desc_sw128(sK, 16, 1024);
    const uint64_t dV = make_smem_desc_sw128(sV, kKVAtom, 1024);  // MN-major: atoms kKVAtom apart
    const uint64_t dP = make_smem_desc_sw128(sP + 2 * t * kQAtom, 16, 1024);
    const uint32_t q_lo = static_cast<uint32_t>(dQ), q_hi = static_cast<uint32_t>(dQ >> 32);
    const uint32_t k_lo = static_cast<uint32_t>(dK), k_hi = static_cast<uint32_t>(dK >> 32);
    const uint32_t v_lo = static_cast<uint32_t>(dV), v_hi = static_cast<uint32_t>(dV >> 32);
    const uint32_t p_lo = static_cast<uint32_t>(dP), p_hi = static_cast<uint32_t>(dP >> 32);
    const uint32_t kv_u = kv_bytes >> 4;
    const uint32_t tS0 = tmem_base + 2 * t * 64;
    const uint32_t tOt = tmem_base + p.o_col + 2 * t * p.o_stride;
    int kst = 0, vst = 0;
    uint32_t kph = 0, vph = 0;
    int cw0 = 0, cw1 = 0;  // completions so far of this tile's per-parity barriers (S full / free, P full, PV done);
                           // two scalars, not an array: a dynamically indexed array lives in local memory
    int n = 0, n_act = 0;  // items seen / items in which this query tile existed
    // S_t into buffer bfs from the current K stage (MMAs only; the commits follow separately)
    auto issue_s = [&](int bfs) {
      const uint32_t b0 = k_lo + kst * kv_u;
      for (int s = 0; s < ksteps_s; ++s) {
        const uint32_t in_atom = (s & 3) * 2;
        umma_f16_ss_w(leader, tS0 + bfs * 64, q_lo + (s >> 2) * (kQAtom >> 4) + in_atom, q_hi,
                      b0 + (s >> 2) * (kKVAtom >> 4) + in_atom, k_hi, idesc_s, s != 0);
      }
    };
    auto commit_s = [&](int bfs) {
      umma_commit_w(leader, (bar0 + 8 * (kBarSFull + 2 * (t) + (bfs))));
      umma_commit_w(leader, (bar0 + 8 * (kBarKEmpty + (kst))));
      if (++kst == p.stages) {
        kst = 0;
        kph ^= 1u;
      }
    };
    // S of tile jj of the current item: its buffer's previous use (this item or the one before) must have been read
    auto s_tile = [&](int jj) {
      const int bfs = jj & 1;
      const int gi = (bfs ? cw1 : cw0) + (jj >> 1);  // how many times this buffer has been used before
      if (gi > 0) mbar_wait_spin((bar0 + 8 * (kBarSFree + 2 * (t) + (bfs))), (gi - 1) & 1);
      mbar_wait_spin((bar0 + 8 * (kBarKFull + (kst))), kph);
      tc_fence_after();
      issue_s(bfs);
      commit_s(bfs);
    };
    for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n) {
      int q0, head, b, ntiles, T;
      item_geom(item, q0, head, b, ntiles, T);
      mbar_wait_spin((bar0 + 8 * kBarQFull), n & 1);
      if (t >= ntiles) {
        // my query tile does not exist in this item (ragged last block): keep the shared barriers in step
        umma_commit_w(leader, (bar0 + 8 * kBarQFree));
        for (int j = 0; j < T; ++j) {
          mbar_wait_spin((bar0 + 8 * (kBarKFull + (kst))), kph);
          umma_commit_w(leader, (bar0 + 8 * (kBarKEmpty + (kst))));
          if (++kst == p.stages) {
            kst = 0;
            kph ^= 1u;
          }
          mbar_wait_spin((bar0 + 8 * (kBarVFull + (vst))), vph);
          umma_commit_w(leader, (bar0 + 8 * (kBarVEmpty + (vst))));
          if (++vst == p.stages) {
            vst = 0;
            vph ^= 1u;
          }
        }
        continue;
      }
      for (int jj = 0; jj < 2 && jj < T; ++jj) s_tile(jj);  // prologue: both S buffers
      if (T <= 2) umma_commit_w(leader, (bar0 + 8 * kBarQFree));  // every S of the item has been issued
      for (int j = 0; j < T; ++j) {
        const int bf = j & 1;
        const int gi = (bf ? cw1 : cw0) + (j >> 1);
        if (j + 2 < T) {
          s_tile(j + 2);
          if (j + 3 >= T) umma_commit_w(leader, (bar0 + 8 * kBarQFree));  // that was the item's last S
        }
        mbar_wait_spin((bar0 + 8 * (kBarPFull + 2 * (t) + (bf))), gi & 1);  // P_t(j) is ready
        mbar_wait_spin((bar0 + 8 * (kBarVFull + (vst))), vph);
        // the first product into an accumulator overwrites it: the previous item's merge must have read it
        if (j < 2 && n_act > 0) mbar_wait_spin((bar0 + 8 * (kBarOFree + t)), (n_act - 1) & 1);
        tc_fence_after();
        const uint32_t a0 = p_lo + bf * (kQAtom >> 4);
        const uint32_t b0 = v_lo + vst * kv_u;
        const uint32_t tOa = tOt + bf * p.o_stride;  // accumulator of the warpgroup that owns tiles of parity bf
        if (p.p_col >= 0) {
          const uint32_t tP = tmem_base + p.p_col + 32 * t;  // 16 keys = 8 columns of 16-bit pairs per step
#pragma unroll
          for (int s = 0; s < kBKV / 16; ++s)
            umma_f16_ts_w(leader, tOa, tP + s * 8, b0 + s * 128, v_hi, idesc_o, (j >= 2) || (s != 0));
        } else {
#pragma unroll
          for (int s = 0; s < kBKV / 16; ++s)  // 16 kv rows per step: +32 B along P's K, +2048 B down V's rows
            umma_f16_ss_w(leader, tOa, a0 + s * 2, p_hi, b0 + s * 128, v_hi, idesc_o, (j >= 2) || (s != 0));
        }
        umma_commit_w(leader, (bar0 + 8 * (kBarVEmpty + (vst))));
        umma_commit_w(leader, (bar0 + 8 * (kBarPVDone + 2 * (t) + (bf))));
        if (++vst == p.stages) {
          vst = 0;
          vph ^= 1u;
        }
      }
      cw0 += (T + 1) >> 1;
      cw1 += T >> 1;
      ++n_act;
    }
  }
  } else {
    // ------------------------------------------------------------------ softmax / correction / epilogue
    pdl_wait();  // the output rows may still be read by an earlier kernel
    const int g = (warp - 4) >> 2;  // warpgroup 0..3
    const int t = g >> 1;           // query tile
    const int wg = g & 1;           // parity of the KV tiles this warpgroup owns (and its S / P buffer, O accumulator)
    const int qd = warp & 3;
    const int r = qd * 32 + lane;  // row in tile == TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
    const uint32_t tS = tmem_base + (2 * t + wg) * 64 + lane_off;
    const uint32_t tO0 = tmem_base + p.o_col + (2 * t) * p.o_stride + lane_off;  // even-tile accumulator
    const uint32_t tO = tO0 + wg * p.o_stride;                                   // mine
    const uint32_t sPt = sP + (2 * t + wg) * kQAtom + r * 128;
    const uint32_t swz = static_cast<uint32_t>(r & 7) << 4;
    const int pair_bar = 1 + t * 4 + qd;  // named barrier shared by the two warps that own the same 32 rows
    int cw0 = 0, cw1 = 0;  // completions so far of this tile's per-parity barriers (as in the issuer)
    for (int item = blockIdx.x; t < p.qtiles && item < p.items; item += gridDim.x) {
      int q0, head, b, ntiles, T;
      item_geom(item, q0, head, b, ntiles, T);
      if (t >= ntiles) continue;
      const int qrow = q0 + t * kBQ + r;
      float mb = 0.f;     // my reference exponent (log2 domain): p = 2^(s * scale_log2 - mb)
      float l_run = 0.f;  // my row sum
      for (int j = wg; j < T; j += 2) {
        const int gi = (wg ? cw1 : cw0) + (j >> 1);  // running index of this use of my S / P barriers
        const int col0 = j * kBKV;
        int limit = p.nk - col0;  // valid columns: [0, limit)
        if (p.causal) limit = min(limit, qrow - col0 + 1);
        const bool need_mask = __any_sync(0xffffffffu, limit < kBKV);
        mbar_wait_spin((bar0 + 8 * (kBarSFull + 2 * (t) + (wg))), gi & 1);
        tc_fence_after();
        // Two passes over the 64 scores in TMEM, 32 at a time through ONE register array: the row maximum first, the
        // exponentials after the reference is settled.  Holding all 64 scores (as an earlier version did, to release S
        // before the exponentials) left no room for the packed row-sum adds and cost a register move per pair; the
        // second read of S is cheap, and S is handed back as soon as its last columns are in registers.
        uint32_t sc[32];
        float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          tmem_ld32(tS + 32 * half, sc);
          tmem_ld_wait();
          if (need_mask) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (i + 32 * half >= limit) sc[i] = __float_as_uint(-INFINITY);
          }
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            m0 = fmaxf(m0, __uint_as_float(sc[i]));
            m1 = fmaxf(m1, __uint_as_float(sc[i + 1]));
            m2 = fmaxf(m2, __uint_as_float(sc[i + 2]));
            m3 = fmaxf(m3, __uint_as_float(sc[i + 3]));
          }
        }
        // Lazily updated reference (0 at the start): it only moves - with a rescale of O / l - when the tile max
        // overshoots it by more than 2^kTau (or, on my first tile, undershoots it by as much); p <= 2^kTau stays far
        // inside fp16 range and the sums are fp32.
        constexpr float kTau = 8.f;
        const float xmax = fmaf(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)), p.scale_log2, -mb);  // -inf: row fully masked
        const bool exceed = (xmax > kTau) || (j == wg && xmax < -kTau && xmax != -INFINITY);
        const uint32_t pv_bar = bar0 + 8 * (kBarPVDone + 2 * t + wg);
        const uint32_t pv_par = (gi - 1) & 1;  // my previous tile (always of this item where it is waited for)
        if (__any_sync(0xffffffffu, exceed)) {  // rare
          const float delta = exceed ? xmax : 0.f;
          const float alpha = ex2f(-delta);
          mb += delta;
          if (j >= 2) {
            l_run *= alpha;  // (on my first tile there is nothing to rescale, and 2^-delta may overflow)
            mbar_wait(pv_bar, pv_par);  // O holds all my earlier tiles
            tc_fence_after();
            for (int c = 0; c < p.dN; c += 16) {
              uint32_t o[16];
              tmem_ld16(tO + c, o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st16(tO + c, o);
            }
            tmem_st_wait();
          }
        }
        // 64 scores -> exp2 -> 16-bit P (registers); row sum by packed fp32x2 adds into two pair accumulators
        float ps[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
        uint32_t w[32];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          tmem_ld32(tS + 32 * half, sc);
          tmem_ld_wait();
          if (half == 1) {  // S is in registers: Q K^T of my next tile may overwrite it
            tc_fence_before();
            mbar_arrive((bar0 + 8 * (kBarSFree + 2 * (t) + (wg))));
          }
          if (need_mask) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (i + 32 * half >= limit) sc[i] = __float_as_uint(-INFINITY);
          }
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            // both scores of the pair through one packed fp32x2 FMA (half the FMA-pipe issue slots)
            float x0, x1;
            fma2(__uint_as_float(sc[2 * k]), __uint_as_float(sc[2 * k + 1]), p.scale_log2, -mb, x0, x1);
            const float p0 = ex2f(x0), p1 = ex2f(x1);
            fadd2(ps[k & 1][0], ps[k & 1][1], p0, p1, ps[k & 1][0], ps[k & 1][1]);
            w[16 * half + k] = pack_act2(p0, p1);
          }
        }
        l_run += (ps[0][0] + ps[1][0]) + (ps[0][1] + ps[1][1]);
        if (p.p_col >= 0) {
          // P through tensor memory: my row's 64 probabilities as 32 packed columns of the query tile's P buffer, which
          // the two warpgroups of the tile share - tile j may be written once P(j-1) V(j-1) (the other warpgroup's
          // tile) has been consumed.  No shared-memory store, and P V reads its A operand from TMEM: half of this
          // kernel's shared-memory traffic is gone.
          if (j >= 1) mbar_wait_spin(bar0 + 8 * (kBarPVDone + 2 * t + (wg ^ 1)), ((wg ? cw0 : cw1) + ((j - 1) >> 1)) & 1);
          tmem_st32(tmem_base + p.p_col + 32 * t + lane_off, w);
          tmem_st_wait();
        } else {
          // P(j-2) V(j-2) has read my P buffer (issued a whole tile ago - the wait is normally already satisfied)
          if (j >= 2) mbar_wait_spin(pv_bar, pv_par);
#pragma unroll
          for (int q = 0; q < 8; ++q)  // 8 swizzled 16-byte chunks of my row
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sPt + ((static_cast<uint32_t>(q) << 4) ^ swz)),
                         "r"(w[q * 4]), "r"(w[q * 4 + 1]), "r"(w[q * 4 + 2]), "r"(w[q * 4 + 3])
                         : "memory");
          fence_proxy_async_smem();
        }
        tc_fence_before();
        mbar_arrive((bar0 + 8 * (kBarPFull + 2 * (t) + (wg))));
      }
      // ---- merge of the two accumulators of my query tile and store.  O = (f_e O_e + f_o O_o) / (f_e l_e + f_o l_o)
      // with f_x = 2^(mb_x - max(mb_e, mb_o)); both accumulators live in my TMEM lanes, only mb / l travel through smem.
      const int mine = (T - wg + 1) >> 1;  // tiles I processed
      const bool has_odd = T >= 2;
      if (mine > 0) mbar_wait((bar0 + 8 * (kBarPVDone + 2 * (t) + (wg))), ((wg ? cw1 : cw0) + mine - 1) & 1);  // the last commit covers every earlier MMA
      const uint32_t xa = sX + ((t * 2 + wg) * 2 * kBQ + r) * 4, xb = sX + ((t * 2 + (wg ^ 1)) * 2 * kBQ + r) * 4;
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(xa), "f"(mb) : "memory");
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(xa + kBQ * 4), "f"(l_run) : "memory");
      tc_fence_before();
      named_bar_sync(pair_bar, 64);
      tc_fence_after();
      float mb_o, l_o;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(mb_o) : "r"(xb) : "memory");
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(l_o) : "r"(xb + kBQ * 4) : "memory");
      named_bar_sync(pair_bar, 64);  // (my slots are rewritten at the end of the NEXT item: the partner has read them)
      // (e, o) = (even-tile warpgroup, odd-tile warpgroup) regardless of which one I am
      const float mb_e = wg ? mb_o : mb, mb_d = wg ? mb : mb_o;
      float l_e = wg ? l_o : l_run, l_d = wg ? l_run : l_o;
      const float mref = has_odd ? fmaxf(mb_e, mb_d) : mb_e;
      const float f_e = ex2f(mb_e - mref), f_d = has_odd ? ex2f(mb_d - mref) : 0.f;
      const float inv_l = 1.f / (f_e * l_e + f_d * l_d);
      const float g_e = f_e * inv_l, g_d = f_d * inv_l;
      act_t* orow = p.out + (static_cast<size_t>(b) * p.nq + qrow) * p.ldo + head * p.d;
      for (int c = wg * 16; c < p.dN; c += 32) {  // the two warpgroups take alternate 16-column chunks
        uint32_t oe[16], od[16];
        tmem_ld16(tO0 + c, oe);
        if (has_odd) tmem_ld16(tO0 + p.o_stride + c, od);
        tmem_ld_wait();
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          v[i] = __uint_as_float(oe[i]) * g_e;
          if (has_odd) v[i] = fmaf(__uint_as_float(od[i]), g_d, v[i]);
        }
        if (qrow < p.nq) {
#pragma unroll
          for (int h8 = 0; h8 < 2; ++h8) {
            if (c + h8 * 8 < p.d) {  // d is a multiple of 8
              uint4 u;
              u.x = pack_act2(v[h8 * 8 + 0], v[h8 * 8 + 1]);
              u.y = pack_act2(v[h8 * 8 + 2], v[h8 * 8 + 3]);
              u.z = pack_act2(v[h8 * 8 + 4], v[h8 * 8 + 5]);
              u.w = pack_act2(v[h8 * 8 + 6], v[h8 * 8 + 7]);
              *reinterpret_cast<uint4*>(orow + c + h8 * 8) = u;
            }
          }
        }
      }
      // O has been read out: the issuer may overwrite the accumulators with the next item's first products
      tc_fence_before();
      mbar_arrive((bar0 + 8 * (kBarOFree + t)));
      cw0 += (T + 1) >> 1;
      cw1 += T >> 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

struct AttnLaunch {
  AttnDev d;
  dim3 grid;
  int smem_bytes;
};

}  // namespace b200sd

using namespace b200sd;

extern "C" int b200sd_attention(const b200sd_attn_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "attention: params is NULL");
  const b200sd_attn_params& h = *hp;
  DeviceInfo di;
  if (int e = get_device_info(&di)) return e;
  if (int e = ensure_dyn_smem(reinterpret_cast<const void*>(&attention_kernel), di.max_smem_optin - 512)) return e;
  {
    AttnLaunch hit;  // identical parameter block seen before: tensor maps and geometry are cached
    if (launch_cache_get(1, hp, sizeof(*hp), &hit, sizeof(hit))) {
      B200SD_CUDA_CHECK(launch_pdl(kPdlAttention, attention_kernel, dim3(hit.grid), dim3(kAttThreads), hit.smem_bytes,
                                   static_cast<cudaStream_t>(stream_), hit.d));
      return B200SD_OK;
    }
  }
  B200SD_REQUIRE(h.q && h.k && h.v && h.out, B200SD_ERR_BAD_ARGUMENT, "attention: NULL pointer");
  B200SD_REQUIRE(h.batch > 0 && h.heads > 0 && h.nq > 0 && h.nk > 0 && h.d > 0, B200SD_ERR_BAD_SHAPE,
                 "attention: non-positive dimension");
  B200SD_REQUIRE((h.d % 8) == 0 && h.d <= 160, B200SD_ERR_BAD_SHAPE,
                 "attention: head dim %d unsupported (multiple of 8, <= 160)", h.d);
  B200SD_REQUIRE((h.ldq % 8) == 0 && (h.ldk % 8) == 0 && (h.ldv % 8) == 0 && (h.ldo % 8) == 0, B200SD_ERR_BAD_SHAPE,
                 "attention: row strides must be multiples of 8 elements");
  B200SD_REQUIRE(h.ldq >= h.heads * h.d && h.ldk >= h.heads * h.d && h.ldv >= h.heads * h.d && h.ldo >= h.heads * h.d,
                 B200SD_ERR_BAD_SHAPE, "attention: row stride smaller than heads*d");
  B200SD_REQUIRE(aligned16(h.q) && aligned16(h.k) && aligned16(h.v) && aligned16(h.out), B200SD_ERR_BAD_ALIGNMENT,
                 "attention: pointers must be 16-byte aligned");
  B200SD_REQUIRE(static_cast<long long>(h.heads) * h.batch * ceil_div(h.nq, kBQ) < (1ll << 30), B200SD_ERR_TOO_LARGE,
                 "attention: too many work items");
  B200SD_REQUIRE(h.scale > 0.f, B200SD_ERR_BAD_ARGUMENT, "attention: scale must be positive");
  if (!h.causal && h.nk <= 128) {
    // the whole key set is one tile (cross-attention against the 77 text tokens): persistent short-KV kernel
    const char* sw = getenv("B200SD_XATTN");  // "0": general kernel (A/B measurements)
    if (sw == nullptr || sw[0] != '0') {
      const int rc = launch_cross_attention(h, di, static_cast<cudaStream_t>(stream_));
      if (rc <= 0) return rc;
    }
  }

  AttnDev d;
  memset(&d, 0, sizeof(d));
  d.out = static_cast<act_t*>(h.out);
  d.batch = h.batch;
  d.heads = h.heads;
  d.nq = h.nq;
  d.nk = h.nk;
  d.d = h.d;
  d.ldo = h.ldo;
  d.nd = ceil_div(h.d, 64);
  d.dN = ceil_div(h.d, 16) * 16;
  d.qtiles = (d.dN <= 64 && h.nq > kBQ) ? 2 : 1;
  // TMEM: S double-buffered per query tile (64 columns each), then the O accumulators
  d.o_col = d.qtiles * 128;
  d.o_stride = d.qtiles == 2 ? 64 : 192;  // 2 x (128 + 2 * 64) or 128 + 2 * 192 = 512 columns
  // P through TMEM when 32 columns per query tile are left beside S and the accumulators (every head dim but 49..64)
  d.p_col = -1;
  {
    const char* sw = getenv("B200SD_ATT_PTMEM");  // "0": P through shared memory (A/B measurements)
    const int ostr = d.qtiles == 2 ? 48 : 160;
    if ((sw == nullptr || sw[0] != '0') && d.dN <= ostr) {
      d.o_stride = ostr;
      d.p_col = d.o_col + 2 * d.qtiles * d.o_stride;  // 256 + 192 = 448 (+ 2 x 32) or 128 + 320 = 448 (+ 32)
    }
  }
  d.causal = h.causal;
  d.scale_log2 = h.scale * 1.4426950408889634f;
  // shared memory: Q + 2 P buffers per query tile are fixed; K/V rings take what is left (at most 4 stages)
  const int fixed = d.qtiles * (d.nd * static_cast<int>(kQAtom) + (d.p_col >= 0 ? 0 : 2 * static_cast<int>(kQAtom)) +
                                2 * 2 * kBQ * 4) + 1024;
  const int per_stage = 2 * d.nd * static_cast<int>(kKVAtom);
  d.stages = min(kMaxStages, (di.max_smem_optin - 1024 - fixed) / per_stage);
  B200SD_REQUIRE(d.stages >= 2, B200SD_ERR_TOO_LARGE, "attention: shared memory budget");

  auto enc = [&](CUtensorMap* tm, const void* ptr, int rows, int ld, uint32_t box_rows) -> int {
    const uint64_t dims[4] = {static_cast<uint64_t>(h.d), static_cast<uint64_t>(h.heads), static_cast<uint64_t>(rows),
                              static_cast<uint64_t>(h.batch)};
    const uint64_t st[3] = {static_cast<uint64_t>(h.d) * 2, static_cast<uint64_t>(ld) * 2,
                            static_cast<uint64_t>(ld) * 2 * rows};
    const uint32_t box[4] = {64, 1, box_rows, 1};
    return encode_tmap_act(tm, ptr, 4, dims, st, box);
  };
  if (int e = enc(&d.tmQ, h.q, h.nq, h.ldq, kBQ)) return e;
  if (int e = enc(&d.tmK, h.k, h.nk, h.ldk, kBKV)) return e;
  if (int e = enc(&d.tmV, h.v, h.nk, h.ldv, kBKV)) return e;

  const int smem_bytes = fixed + d.stages * per_stage;
  d.nqblk = ceil_div(h.nq, kBQ * d.qtiles);
  d.items = h.batch * h.heads * d.nqblk;
  dim3 grid(min(d.items, di.sm_count));  // persistent: one CTA per SM, items taken with stride gridDim.x
  if (const char* sw = getenv("B200SD_ATT_PERSIST"))  // "0": one CTA per item (A/B measurements)
    if (sw[0] == '0') grid = dim3(d.items);
  {
    AttnLaunch L;
    memset(&L, 0, sizeof(L));
    L.d = d;
    L.grid = grid;
    L.smem_bytes = smem_bytes;
    launch_cache_put(1, hp, sizeof(*hp), &L, sizeof(L));
  }
  B200SD_CUDA_CHECK(launch_pdl(kPdlAttention, attention_kernel, dim3(grid), dim3(kAttThreads), smem_bytes,
                               static_cast<cudaStream_t>(stream_), d));
  return B200SD_OK;
}
