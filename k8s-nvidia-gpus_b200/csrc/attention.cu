// Flash attention forward on tcgen05 (sm_100a): out = softmax(Q K^T * scale [+ causal]) V per (batch, head).
// SURVEY §2.1 K5 (self-attention), K6 (cross-attention, nk = 77), K15 (CLIP causal attention).
//
// One CTA per (128-query tile, head, batch):
//   warp 0 (1 thread) : TMA producer.  Q/K/V are read straight out of the projection GEMM outputs through 4-D
//                       tensor maps {d, head, token, batch}; the box is 64 wide in d, so head dims that are not a
//                       multiple of 64 (40, 80, 160) and ragged token counts are zero-filled by the TMA unit —
//                       no padded copies of Q/K/V exist in HBM.
//   warp 1 (1 thread) : tcgen05.mma issuer.  S = Q K^T (K-major A and B) into TMEM columns [0,128);
//                       O += P V with P (act_t, written by the softmax warps) as the K-major A operand in shared
//                       memory and V consumed in place as an MN-major B operand; O lives in TMEM columns [128, ...).
//   warp 2            : TMEM allocator.
//   warps 4..7        : online softmax (thread = query row): tcgen05.ld S, running max / sum in fp32 registers,
//                       exp2 with the scale folded in, P -> swizzled shared memory, lazy rescale of O in TMEM,
//                       final normalisation and act_t store.
// S_{j+1} is issued before P_j V_j so the tensor pipe computes the next scores while the softmax warps work.
#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kAttThreads = 256;
static constexpr int kBQ = 128;   // query rows per CTA (UMMA M)
static constexpr int kBKV = 128;  // keys per tile
static constexpr uint32_t kAtomBytes = 128 * 128;  // 128 rows x 64 act_t, one swizzle-128B atom column

struct alignas(64) AttnDev {
  CUtensorMap tmQ, tmK, tmV;
  act_t* out;
  int batch, heads, nq, nk, d;
  int ldo;
  int nd;      // 64-wide d atoms
  int dN;      // d rounded up to 16 (UMMA N of the PV product, K extent of QK^T)
  int stages;  // K/V ring depth (1 or 2)
  int tmem_cols;
  int causal;
  float scale_log2;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAttThreads) attention_kernel(const __grid_constant__ AttnDev p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full, s_full, p_full, pv_done;
  __shared__ __align__(8) uint64_t k_full[2], v_full[2], k_empty[2], v_empty[2];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kBQ;
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t tile_bytes = p.nd * kAtomBytes;
  const uint32_t sQ = base;
  const uint32_t sK = sQ + tile_bytes;
  const uint32_t sV = sK + p.stages * tile_bytes;
  const uint32_t sP = sV + p.stages * tile_bytes;
  int kv_end = p.nk;
  if (p.causal) kv_end = min(p.nk, q0 + kBQ);  // keys beyond the last query row of this tile are fully masked
  const int T = (kv_end + kBKV - 1) / kBKV;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQ);
    tma_prefetch_desc(&p.tmK);
    tma_prefetch_desc(&p.tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(smem_u32(&q_full), 1);
    mbar_init(smem_u32(&s_full), 1);
    mbar_init(smem_u32(&p_full), 128);
    mbar_init(smem_u32(&pv_done), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&k_full[s]), 1);
      mbar_init(smem_u32(&v_full[s]), 1);
      mbar_init(smem_u32(&k_empty[s]), 1);
      mbar_init(smem_u32(&v_empty[s]), 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(smem_u32(&tmem_base_s), p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  const uint32_t tS = tmem_base;        // 128 fp32 columns
  const uint32_t tO = tmem_base + 128;  // dN fp32 columns

  if (warp == 0) {
    if (lane == 0) {
      // ---------------------------------------------------------------- TMA producer
      mbar_arrive_expect_tx(smem_u32(&q_full), tile_bytes);
      for (int a = 0; a < p.nd; ++a) tma_load_4d(sQ + a * kAtomBytes, &p.tmQ, smem_u32(&q_full), a * 64, head, q0, b);
      for (int j = 0; j < T; ++j) {
        const int st = j % p.stages;
        const uint32_t ph = ((j / p.stages) & 1) ^ 1u;
        mbar_wait(smem_u32(&k_empty[st]), ph);
        mbar_arrive_expect_tx(smem_u32(&k_full[st]), tile_bytes);
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d(sK + st * tile_bytes + a * kAtomBytes, &p.tmK, smem_u32(&k_full[st]), a * 64, head, j * kBKV, b);
        mbar_wait(smem_u32(&v_empty[st]), ph);
        mbar_arrive_expect_tx(smem_u32(&v_full[st]), tile_bytes);
        for (int a = 0; a < p.nd; ++a)
          tma_load_4d(sV + st * tile_bytes + a * kAtomBytes, &p.tmV, smem_u32(&v_full[st]), a * 64, head, j * kBKV, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ---------------------------------------------------------------- MMA issuer
      const uint32_t idesc_s = make_idesc_f16(kBQ, kBKV, 0, 0);  // S[128 x 128] = Q (K-major) x K (K-major)
      const uint32_t idesc_o = make_idesc_f16(kBQ, p.dN, 0, 1);  // O[128 x dN] += P (K-major) x V (MN-major)
      const int ksteps_s = p.dN / 16;
      auto issue_s = [&](int j) {
        const int st = j % p.stages;
        mbar_wait(smem_u32(&k_full[st]), (j / p.stages) & 1);
        tc_fence_after();
        for (int s = 0; s < ksteps_s; ++s) {
          const uint32_t off = (s >> 2) * kAtomBytes + (s & 3) * 32;
          const uint64_t adesc = make_smem_desc_sw128(sQ + off, 16, 1024);
          const uint64_t bdesc = make_smem_desc_sw128(sK + st * tile_bytes + off, 16, 1024);
          umma_f16_ss(tS, adesc, bdesc, idesc_s, s != 0);
        }
        umma_commit(smem_u32(&k_empty[st]));
        umma_commit(smem_u32(&s_full));
      };
      mbar_wait(smem_u32(&q_full), 0);
      issue_s(0);
      for (int j = 0; j < T; ++j) {
        const int st = j % p.stages;
        mbar_wait(smem_u32(&p_full), j & 1);  // S_j consumed, P_j in smem, O rescaled
        tc_fence_after();
        if (j + 1 < T) issue_s(j + 1);
        mbar_wait(smem_u32(&v_full[st]), (j / p.stages) & 1);
        tc_fence_after();
#pragma unroll
        for (int s = 0; s < kBKV / 16; ++s) {
          const uint64_t adesc = make_smem_desc_sw128(sP + (s >> 2) * kAtomBytes + (s & 3) * 32, 16, 1024);
          // V tile: [kv rows at 128 B pitch][64 d per atom]; 16 kv rows per step = 2048 B; atoms kAtomBytes apart
          const uint64_t bdesc = make_smem_desc_sw128(sV + st * tile_bytes + s * 2048, kAtomBytes, 1024);
          umma_f16_ss(tO, adesc, bdesc, idesc_o, (j | s) != 0);
        }
        umma_commit(smem_u32(&v_empty[st]));
        umma_commit(smem_u32(&pv_done));
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ softmax / correction / epilogue
    const int qd = warp & 3;
    const int r = qd * 32 + lane;  // row in tile == TMEM lane
    const int qrow = q0 + r;
    const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
    float m_run = -INFINITY;
    float l_run = 0.f;
    for (int j = 0; j < T; ++j) {
      mbar_wait(smem_u32(&s_full), j & 1);
      tc_fence_after();
      uint32_t sv[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tS + lane_off + c * 32, sv[c]);
      tmem_ld_wait();
      // mask + tile max
      const int col0 = j * kBKV;
      int limit = p.nk - col0;                                   // valid columns: [0, limit)
      if (p.causal) limit = min(limit, qrow - col0 + 1);
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = __uint_as_float(sv[c][i]);
          s = (c * 32 + i < limit) ? s : -INFINITY;
          sv[c][i] = __float_as_uint(s);
          mx = fmaxf(mx, s);
        }
      const float m_new = fmaxf(m_run, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = ex2f((m_run - m_use) * p.scale_log2);  // m_run = -inf -> 0
      const float mb = m_use * p.scale_log2;
      float psum = 0.f;
      // P_{j-1} V_{j-1} must have completed before P smem is overwritten / O is rescaled
      if (j > 0) {
        mbar_wait(smem_u32(&pv_done), (j - 1) & 1);
        tc_fence_after();
      }
#pragma unroll
      for (int c16 = 0; c16 < 16; ++c16) {
        float pv[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float s = __uint_as_float(sv[c16 >> 2][(c16 & 3) * 8 + e]);
          pv[e] = ex2f(fmaf(s, p.scale_log2, -mb));
          psum += pv[e];
        }
        const uint32_t addr = sP + (c16 >> 3) * kAtomBytes + r * 128 + (((c16 & 7) ^ (r & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pack_act2(pv[0], pv[1])),
                     "r"(pack_act2(pv[2], pv[3])), "r"(pack_act2(pv[4], pv[5])), "r"(pack_act2(pv[6], pv[7]))
                     : "memory");
      }
      l_run = l_run * alpha + psum;
      m_run = m_new;
      if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
        for (int c = 0; c < p.dN; c += 16) {
          uint32_t o[16];
          tmem_ld16(tO + lane_off + c, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st16(tO + lane_off + c, o);
        }
        tmem_st_wait();
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&p_full));
    }
    // final: O / l -> act_t -> global
    mbar_wait(smem_u32(&pv_done), (T - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.f / l_run;
    act_t* orow = p.out + (static_cast<size_t>(b) * p.nq + qrow) * p.ldo + head * p.d;
    for (int c = 0; c < p.dN; c += 16) {
      uint32_t o[16];
      tmem_ld16(tO + lane_off + c, o);
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

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
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
  d.stages = (d.nd == 2) ? 2 : 1;
  d.tmem_cols = (128 + d.dN <= 256) ? 256 : 512;
  d.causal = h.causal;
  d.scale_log2 = h.scale * 1.4426950408889634f;

  auto enc = [&](CUtensorMap* tm, const void* ptr, int rows, int ld) -> int {
    const uint64_t dims[4] = {static_cast<uint64_t>(h.d), static_cast<uint64_t>(h.heads), static_cast<uint64_t>(rows),
                              static_cast<uint64_t>(h.batch)};
    const uint64_t st[3] = {static_cast<uint64_t>(h.d) * 2, static_cast<uint64_t>(ld) * 2,
                            static_cast<uint64_t>(ld) * 2 * rows};
    const uint32_t box[4] = {64, 1, 128, 1};
    return encode_tmap_act(tm, ptr, 4, dims, st, box);
  };
  if (int e = enc(&d.tmQ, h.q, h.nq, h.ldq)) return e;
  if (int e = enc(&d.tmK, h.k, h.nk, h.ldk)) return e;
  if (int e = enc(&d.tmV, h.v, h.nk, h.ldv)) return e;

  const int smem_bytes = d.nd * kAtomBytes * (1 + 2 * d.stages) + 2 * kAtomBytes + 1024;
  if (!g_attn_attr_set) {
    B200SD_CUDA_CHECK(
        cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, di.max_smem_optin - 512));
    g_attn_attr_set = true;
  }
  dim3 grid(ceil_div(h.nq, kBQ), h.heads, h.batch);
  attention_kernel<<<grid, kAttThreads, smem_bytes, static_cast<cudaStream_t>(stream_)>>>(d);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}
