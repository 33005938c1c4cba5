// Flash attention forward on tcgen05 (sm_100a): out = softmax(Q K^T * scale [+ causal]) V per (batch, head).
// SURVEY §2.1 K5 (self-attention), K6 (cross-attention, nk = 77), K15 (CLIP causal attention).
//
// One CTA per (up to two 128-query tiles, head, batch); keys are walked in tiles of 64; 12 warps:
//   warp 0 (1 thread) : TMA producer.  Q/K/V are read straight out of the projection GEMM outputs through 4-D
//                       tensor maps {d, head, token, batch}; the box is 64 wide in d, so head dims that are not a
//                       multiple of 64 (40, 80, 160) and ragged token counts are zero-filled by the TMA unit —
//                       no padded copies of Q/K/V exist in HBM.  K/V tiles are shared by both query tiles.
//   warps 1, 3 (1 thread each): tcgen05.mma issuer of query tile 0 / 1.  S_t = Q_t K^T (K-major A and B) into a DOUBLE-BUFFERED TMEM tile per
//                       query tile; O_t += P_t V with P_t (16-bit, written by softmax warpgroup t, double-buffered in
//                       shared memory) as the K-major A operand and V consumed in place as an MN-major B operand.
//                       S_t(j+2) is issued the moment warpgroup t has consumed S_t(j): the softmax warps never wait
//                       for the tensor pipe's round trip, and the tensor pipe never waits for the softmax of the tile
//                       it just produced.
//   warp 2            : TMEM allocator.
//   warps 4..7 / 8..11: softmax warpgroup 0 / 1 (thread = query row).  One pass per KV tile: tcgen05.ld S (next
//                       chunk in flight) -> p = exp2(s * scale - ref) against a lazily updated reference (moved,
//                       with a rescale of O in TMEM, only when the tile max overshoots it by 2^8) -> 16-bit P ->
//                       swizzled shared memory.  Final normalisation and store.
//   Row sums: when the head dim leaves a spare UMMA column (d = 40 -> N = 48) column d of every V tile is set to
//   1.0 after the TMA lands, so the tensor pipe accumulates sum_j P_ij in fp32 as column d of O for free (and it is
//   rescaled together with O); otherwise the sum is accumulated in registers.
#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kAttThreads = 384;
static constexpr int kBQ = 128;  // query rows per tile (UMMA M)
static constexpr int kBKV = 64;  // keys per tile
static constexpr uint32_t kQAtom = 128 * 128;   // Q / P: 128 rows x 64 elements (one swizzle-128B atom column)
static constexpr uint32_t kKVAtom = kBKV * 128;  // K / V: 64 rows x 64 elements
static constexpr int kMaxStages = 4;

struct alignas(64) AttnDev {
  CUtensorMap tmQ, tmK, tmV;
  act_t* out;
  int batch, heads, nq, nk, d;
  int ldo;
  int nd;         // 64-wide d atoms
  int dN;         // UMMA N of the PV product / K extent of QK^T (d rounded up to 16)
  int stages;     // K/V ring depth
  int qtiles;     // query tiles per CTA (1 or 2)
  int o_col;      // first TMEM column of O_0
  int o_stride;   // TMEM column stride between O_0 and O_1
  int sum_col;    // >= 0: column of O that accumulates the row sums (ones column in V); -1: register sums
  int causal;
  float scale_log2;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <bool kOnes>
__global__ void __launch_bounds__(kAttThreads, 1) attention_kernel(const __grid_constant__ AttnDev p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full;
  __shared__ __align__(8) uint64_t s_full[2][2], p_full[2][2], pv_done[2][2];  // [query tile][buffer]
  __shared__ __align__(8) uint64_t k_full[kMaxStages], v_full[kMaxStages], k_empty[kMaxStages], v_empty[kMaxStages];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kBQ * p.qtiles;
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int ntiles = (p.qtiles == 2 && q0 + kBQ < p.nq) ? 2 : 1;  // query tiles this CTA really has
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_bytes = p.nd * kQAtom;
  const uint32_t kv_bytes = p.nd * kKVAtom;
  const uint32_t sQ = base;                          // [qtiles][nd atoms of 16 KB]
  const uint32_t sK = sQ + p.qtiles * q_bytes;       // [stages][nd atoms of 8 KB]
  const uint32_t sV = sK + p.stages * kv_bytes;      // [stages][nd atoms of 8 KB]
  const uint32_t sP = sV + p.stages * kv_bytes;      // [qtiles][2 buffers][16 KB]
  int kv_end = p.nk;
  if (p.causal) kv_end = min(p.nk, q0 + kBQ * ntiles);  // keys beyond the last query row are fully masked
  const int T = (kv_end + kBKV - 1) / kBKV;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQ);
    tma_prefetch_desc(&p.tmK);
    tma_prefetch_desc(&p.tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(smem_u32(&q_full), 1);
    for (int t = 0; t < 2; ++t)
      for (int i = 0; i < 2; ++i) {
        mbar_init(smem_u32(&s_full[t][i]), 1);
        mbar_init(smem_u32(&p_full[t][i]), 128);
        mbar_init(smem_u32(&pv_done[t][i]), 1);
      }
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(smem_u32(&k_full[s]), 1);
      mbar_init(smem_u32(&v_full[s]), 1);
      mbar_init(smem_u32(&k_empty[s]), ntiles);  // one tcgen05.commit per MMA issuer (one issuer per query tile)
      mbar_init(smem_u32(&v_empty[s]), ntiles);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(smem_u32(&tmem_base_s), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  // TMEM columns: S_t[buf] at (2t + buf) * 64; O_t at o_col + t * o_stride

  if (warp == 0) {
    if (lane == 0) {
      // ---------------------------------------------------------------- TMA producer
      mbar_arrive_expect_tx(smem_u32(&q_full), ntiles * q_bytes);
      for (int t = 0; t < ntiles; ++t)
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d(sQ + t * q_bytes + a * kQAtom, &p.tmQ, smem_u32(&q_full), a * 64, head, q0 + t * kBQ, b);
      int st = 0;
      uint32_t ph = 1;
      for (int j = 0; j < T; ++j) {
        mbar_wait(smem_u32(&k_empty[st]), ph);
        mbar_arrive_expect_tx(smem_u32(&k_full[st]), kv_bytes);
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d(sK + st * kv_bytes + a * kKVAtom, &p.tmK, smem_u32(&k_full[st]), a * 64, head, j * kBKV, b);
        mbar_wait(smem_u32(&v_empty[st]), ph);
        mbar_arrive_expect_tx(smem_u32(&v_full[st]), kv_bytes);
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d(sV + st * kv_bytes + a * kKVAtom, &p.tmV, smem_u32(&v_full[st]), a * 64, head, j * kBKV, b);
        if (++st == p.stages) {
          st = 0;
          ph ^= 1u;
        }
      }
    }
  } else if (warp == 1 || warp == 3) {
    const int t = warp >> 1;  // warp 1 issues for query tile 0, warp 3 for query tile 1
    if (lane == 0 && t < ntiles) {
      // ---------------------------------------------------------------- MMA issuer of query tile t
      // Descriptors are built once; the loop only adds 16-byte-unit offsets to their low words.  K and V stages are
      // consumed strictly in order, so each ring is tracked by a running (stage, phase) pair - no divisions.
      const uint32_t idesc_s = make_idesc_f16(kBQ, kBKV, 0, 0);  // S[128 x 64] = Q (K-major) x K (K-major)
      const uint32_t idesc_o = make_idesc_f16(kBQ, p.dN, 0, 1);  // O[128 x dN] += P (K-major) x V (MN-major)
      const int ksteps_s = p.dN / 16;
      const uint64_t dQ = make_smem_desc_sw128(sQ + t * q_bytes, 16, 1024);
      const uint64_t dK = make_smem_desc_sw128(sK, 16, 1024);
      const uint64_t dV = make_smem_desc_sw128(sV, kKVAtom, 1024);  // MN-major: atoms kKVAtom apart
      const uint64_t dP = make_smem_desc_sw128(sP + 2 * t * kQAtom, 16, 1024);
      const uint32_t kv_u = kv_bytes >> 4;
      const uint32_t tS0 = tmem_base + 2 * t * 64;
      const uint32_t tOt = tmem_base + p.o_col + t * p.o_stride;
      int kst = 0, vst = 0;
      uint32_t kph = 0, vph = 0;
      // S_t(jj) into buffer jj & 1 from the current K stage
      auto issue_s = [&](int bfs) {
        const uint64_t b0 = dK + kst * kv_u;
        for (int s = 0; s < ksteps_s; ++s) {
          const uint32_t in_atom = (s & 3) * 2;
          umma_f16_ss(tS0 + bfs * 64, dQ + (s >> 2) * (kQAtom >> 4) + in_atom, b0 + (s >> 2) * (kKVAtom >> 4) + in_atom,
                      idesc_s, s != 0);
        }
        umma_commit(smem_u32(&s_full[t][bfs]));
        umma_commit(smem_u32(&k_empty[kst]));
        if (++kst == p.stages) {
          kst = 0;
          kph ^= 1u;
        }
      };
      mbar_wait_spin(smem_u32(&q_full), 0);
      for (int jj = 0; jj < 2 && jj < T; ++jj) {  // prologue: both S buffers
        mbar_wait_spin(smem_u32(&k_full[kst]), kph);
        tc_fence_after();
        issue_s(jj);
      }
      for (int j = 0; j < T; ++j) {
        const int bf = j & 1;
        mbar_wait_spin(smem_u32(&p_full[t][bf]), (j >> 1) & 1);  // P_t(j) in smem; S_t buffer bf consumed
        tc_fence_after();
        if (j + 2 < T) {
          mbar_wait_spin(smem_u32(&k_full[kst]), kph);
          tc_fence_after();
          issue_s(bf);
        }
        mbar_wait_spin(smem_u32(&v_full[vst]), vph);
        tc_fence_after();
        const uint64_t a0 = dP + bf * (kQAtom >> 4);
        const uint64_t b0 = dV + vst * kv_u;
#pragma unroll
        for (int s = 0; s < kBKV / 16; ++s)  // 16 kv rows per step: +32 B along P's K, +2048 B down V's rows
          umma_f16_ss(tOt, a0 + s * 2, b0 + s * 128, idesc_o, (j | s) != 0);
        umma_commit(smem_u32(&v_empty[vst]));
        umma_commit(smem_u32(&pv_done[t][bf]));
        if (++vst == p.stages) {
          vst = 0;
          vph ^= 1u;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ softmax / correction / epilogue
    const int t = (warp - 4) >> 2;  // warpgroup = query tile
    if (t < ntiles) {
      const int qd = warp & 3;
      const int r = qd * 32 + lane;  // row in tile == TMEM lane
      const int qrow = q0 + t * kBQ + r;
      const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
      const uint32_t tO = tmem_base + p.o_col + t * p.o_stride + lane_off;
      constexpr bool ones_col = kOnes;
      float mb = 0.f;  // reference exponent (log2 domain): p = 2^(s * scale_log2 - mb)
      int vst = 0;     // V ring position (ones-column writer)
      uint32_t vph = 0;
      float l_run = 0.f;
      for (int j = 0; j < T; ++j) {
        const int bf = j & 1;
        const uint32_t tS = tmem_base + (2 * t + bf) * 64 + lane_off;
        const uint32_t sPt = sP + (2 * t + bf) * kQAtom + r * 128;
        const int col0 = j * kBKV;
        int limit = p.nk - col0;  // valid columns: [0, limit)
        if (p.causal) limit = min(limit, qrow - col0 + 1);
        const bool need_mask = __any_sync(0xffffffffu, limit < kBKV);
        mbar_wait_spin(smem_u32(&s_full[t][bf]), (j >> 1) & 1);
        tc_fence_after();
        if (j >= 2) {
          // P_t(j-2) V(j-2) has been consumed: P buffer bf may be overwritten
          mbar_wait_spin(smem_u32(&pv_done[t][bf]), ((j >> 1) - 1) & 1);
        }
        // Single pass per tile: exponentials against the CURRENT reference while the tile's own max is tracked.  The
        // reference (0 at the start) only moves - with a rescale of O / l - when the tile max overshoots it by more
        // than 2^kTau (or, on the first tile, undershoots it by as much); p <= 2^kTau stays far inside fp16 range and
        // the sums are fp32.  The pass is then redone once.
        constexpr float kTau = 8.f;
        for (int attempt = 0; attempt < 2; ++attempt) {
          float psum = 0.f;
          float m0 = -INFINITY, m1 = -INFINITY;
          uint32_t bufA[32], bufB[32];
          // 32 scores -> exp2 -> 16-bit P -> 4 swizzled 16-byte chunks of my row
          auto process = [&](uint32_t(&cur)[32], int c) {
            if (need_mask) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (c * 32 + i >= limit) cur[i] = __float_as_uint(-INFINITY);
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint32_t w[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float s0 = __uint_as_float(cur[g * 8 + 2 * e]), s1 = __uint_as_float(cur[g * 8 + 2 * e + 1]);
                m0 = fmaxf(m0, s0);
                m1 = fmaxf(m1, s1);
                const float p0 = ex2f(fmaf(s0, p.scale_log2, -mb)), p1 = ex2f(fmaf(s1, p.scale_log2, -mb));
                if (!ones_col) psum += p0 + p1;
                w[e] = pack_act2(p0, p1);
              }
              const int c16 = c * 4 + g;
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sPt + ((c16 ^ (r & 7)) << 4)), "r"(w[0]),
                           "r"(w[1]), "r"(w[2]), "r"(w[3])
                           : "memory");
            }
          };
          tmem_ld32(tS, bufA);
          tmem_ld_wait();
          tmem_ld32(tS + 32, bufB);  // in flight during the first half's maths
          process(bufA, 0);
          tmem_ld_wait();
          process(bufB, 1);
          const float xmax = fmaf(fmaxf(m0, m1), p.scale_log2, -mb);  // -inf when the whole row is masked
          const bool exceed = (xmax > kTau) || (j == 0 && xmax < -kTau && xmax != -INFINITY);
          if (!__any_sync(0xffffffffu, exceed)) {
            l_run += psum;
            break;
          }
          // rare: move the reference of the rows that over/undershot, rescale their O / l, redo the pass once
          const float delta = exceed ? xmax : 0.f;
          const float alpha = ex2f(-delta);
          mb += delta;
          l_run *= alpha;
          if (j > 0) {
            // O must hold every product up to tile j-1
            mbar_wait(smem_u32(&pv_done[t][bf ^ 1]), ((j - 1) >> 1) & 1);
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
        if (ones_col) {
          // V(j) has landed: set column `sum_col` of its 64 kv rows to 1.0 so O[:, sum_col] accumulates the row sums
          mbar_wait_spin(smem_u32(&v_full[vst]), vph);
          if (r < kBKV) {
            const int sc = p.sum_col;
            const uint32_t va = sV + vst * kv_bytes + (sc >> 6) * kKVAtom + r * 128 +
                                (((((sc & 63) >> 3)) ^ (r & 7)) << 4) + (sc & 7) * 2;
            const act_t one = to_act(1.f);
            asm volatile("st.shared.b16 [%0], %1;" ::"r"(va), "h"(*reinterpret_cast<const uint16_t*>(&one)) : "memory");
          }
          if (++vst == p.stages) {
            vst = 0;
            vph ^= 1u;
          }
        }
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(smem_u32(&p_full[t][bf]));
      }
      // final: O / l -> 16-bit -> global (the last commit covers every earlier MMA)
      mbar_wait(smem_u32(&pv_done[t][(T - 1) & 1]), ((T - 1) >> 1) & 1);
      tc_fence_after();
      act_t* orow = p.out + (static_cast<size_t>(b) * p.nq + qrow) * p.ldo + head * p.d;
      float inv_l = 1.f / l_run;
      if (ones_col) {
        uint32_t o[16];
        tmem_ld16(tO + (p.sum_col & ~15), o);
        tmem_ld_wait();
        float l = 0.f;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (i == (p.sum_col & 15)) l = __uint_as_float(o[i]);
        inv_l = 1.f / l;
      }
      for (int c = 0; c < p.dN; c += 16) {
        uint32_t o[16];
        tmem_ld16(tO + c, o);
        tmem_ld_wait();
        if (qrow < p.nq) {
#pragma unroll
          for (int h8 = 0; h8 < 2; ++h8) {
            if (c + h8 * 8 < p.d) {  // d is a multiple of 8
              uint4 u;
              u.x = pack_act2(__uint_as_float(o[h8 * 8 + 0]) * inv_l, __uint_as_float(o[h8 * 8 + 1]) * inv_l);
              u.y = pack_act2(__uint_as_float(o[h8 * 8 + 2]) * inv_l, __uint_as_float(o[h8 * 8 + 3]) * inv_l);
              u.z = pack_act2(__uint_as_float(o[h8 * 8 + 4]) * inv_l, __uint_as_float(o[h8 * 8 + 5]) * inv_l);
              u.w = pack_act2(__uint_as_float(o[h8 * 8 + 6]) * inv_l, __uint_as_float(o[h8 * 8 + 7]) * inv_l);
              *reinterpret_cast<uint4*>(orow + c + h8 * 8) = u;
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

static bool g_attn_attr_set = false;

}  // namespace b200sd

using namespace b200sd;

extern "C" int b200sd_attention(const b200sd_attn_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "attention: params is NULL");
  const b200sd_attn_params& h = *hp;
  DeviceInfo di;
  if (int e = get_device_info(&di)) return e;
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
  B200SD_REQUIRE(h.heads <= 65535 && h.batch <= 65535, B200SD_ERR_TOO_LARGE, "attention: grid too large");
  B200SD_REQUIRE(h.scale > 0.f, B200SD_ERR_BAD_ARGUMENT, "attention: scale must be positive");

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
  d.qtiles = (d.dN <= 128 && h.nq > kBQ) ? 2 : 1;
  // TMEM: S double-buffered per query tile (64 columns each), then the O accumulators
  d.o_col = d.qtiles * 128;
  d.o_stride = 128;
  d.sum_col = (d.dN > h.d) ? h.d : -1;
  d.causal = h.causal;
  d.scale_log2 = h.scale * 1.4426950408889634f;
  // shared memory: Q + 2 P buffers per query tile are fixed; K/V rings take what is left (at most 4 stages)
  const int fixed = d.qtiles * (d.nd * static_cast<int>(kQAtom) + 2 * static_cast<int>(kQAtom)) + 1024;
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
  if (!g_attn_attr_set) {
    B200SD_CUDA_CHECK(cudaFuncSetAttribute(attention_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           di.max_smem_optin - 512));
    B200SD_CUDA_CHECK(cudaFuncSetAttribute(attention_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           di.max_smem_optin - 512));
    g_attn_attr_set = true;
  }
  dim3 grid(ceil_div(h.nq, kBQ * d.qtiles), h.heads, h.batch);
  if (d.sum_col >= 0)
    attention_kernel<true><<<grid, kAttThreads, smem_bytes, static_cast<cudaStream_t>(stream_)>>>(d);
  else
    attention_kernel<false><<<grid, kAttThreads, smem_bytes, static_cast<cudaStream_t>(stream_)>>>(d);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}
