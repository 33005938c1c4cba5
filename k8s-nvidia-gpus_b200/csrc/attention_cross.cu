// Short-KV attention on tcgen05 (sm_100a): out = softmax(Q K^T * scale) V for nk <= 128 keys - the cross-attention of
// BasicTransformerBlock.attn2 against the 77 text tokens.  16 of the 32 attention launches of a
// denoise step; in the general flash kernel (attention.cu) each of them is a chain of latencies with nothing to
// overlap (one CTA = two query tiles = launch, TMEM alloc, Q/K/V loads, two S tiles, softmax, P V, merge, store: 8 us
// of which 0.4 us are arithmetic).  Here the whole key set is ONE tile, so there is no online softmax, no rescale and
// no accumulator merge, and the kernel is persistent:
//
//   grid = SM count; every CTA walks a contiguous range of work items (128 query rows of one (batch, head)).
//   K and V of a (batch, head) are loaded once per run of items (two sets, so the next run's K/V land early).
//   NP independent pipelines (softmax warpgroup + Q / P shared-memory tiles + S / O accumulators in TMEM) take the
//   items round-robin, so the Q load and Q K^T of the next items run under the exponentials and the store of the
//   current one:
//     warp 0      TMA producer (K/V per run, Q per item)
//     warp 1      tcgen05.mma issuer, polling: S_k = Q_k K^T (N = keys rounded to 16) as soon as pipe k % NP is free,
//                 O_k = P_k V as soon as P_k is in shared memory
//     warp 2      TMEM allocator
//     warps 4..   NP softmax warpgroups (thread = query row): S -> registers, row max, p = exp2(s * c - m), row sum,
//                 16-bit P -> swizzled shared memory; then O -> registers, * 1/sum, 16-byte stores
//   Only the key columns that exist are exponentiated (77 -> 80, not 128).
#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kXMaxPipes = 3;
static constexpr int kXBQ = 128;
static constexpr uint32_t kXQAtom = 128 * 128;  // Q / P atom: 128 rows x 64 elements

enum : int {
  kXQFull = 0,                     // [pipe] Q tile landed
  kXQFree = kXQFull + kXMaxPipes,  // [pipe] S MMAs that read it have completed
  kXSFull = kXQFree + kXMaxPipes,  // [pipe] S in TMEM
  kXPFull = kXSFull + kXMaxPipes,  // [pipe] P in shared memory
  kXOFull = kXPFull + kXMaxPipes,  // [pipe] O in TMEM
  kXPipeFree = kXOFull + kXMaxPipes,  // [pipe] O read out, P / S reusable
  kXKVFull = kXPipeFree + kXMaxPipes,  // [set] K and V of a run landed
  kXKVFree = kXKVFull + 2,             // [set] last P V of the run completed
  kXNumBars = kXKVFree + 2
};

struct alignas(64) XAttnDev {
  CUtensorMap tmQ, tmK, tmV;
  act_t* out;
  int batch, heads, nq, nk, d, ldo;
  int nd;       // 64-wide d atoms
  int dN;       // d rounded up to 16 (UMMA N of P V, K extent of Q K^T)
  int KP;       // keys rounded up to 16 (UMMA N of Q K^T, K extent of P V)
  int np;       // pipelines
  int tstride;  // TMEM columns per pipeline (S at +0, O at +KP)
  int qtiles;   // query tiles per (batch, head)
  int items;    // batch * heads * qtiles
  int ipc;      // items per CTA (contiguous)
  float scale_log2;
};

// non-blocking phase test (try_wait may park the thread for a system-dependent time before it answers "not yet" -
// a polling scheduler must not)
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}

__device__ __forceinline__ float x_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// kCh: 16-column score chunks a row holds in registers (5 = up to 80 keys: the 77 text tokens; 8 = up to 128 keys);
// kPipes: most pipelines the launch may use (sets the register budget per thread).
template <int kCh, int kPipes>
__global__ void __launch_bounds__(128 + 128 * kPipes, 1) xattn_kernel(const __grid_constant__ XAttnDev p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[kXNumBars];
  __shared__ uint32_t tmem_base_s;
  const uint32_t bar0 = smem_u32(bars);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_bytes = p.nd * kXQAtom;                 // one Q tile
  const uint32_t kv_atom = static_cast<uint32_t>(p.KP) * 128u;  // K / V atom: KP rows x 64 elements
  const uint32_t kv_bytes = p.nd * kv_atom;                // one K (or V) tile
  const uint32_t p_atoms = static_cast<uint32_t>((p.KP + 63) / 64);
  const uint32_t sQ = base;                                // [np][nd atoms]
  const uint32_t sP = sQ + p.np * q_bytes;                 // [np][p_atoms atoms]
  const uint32_t sK = sP + p.np * p_atoms * kXQAtom;       // [2 sets][nd atoms]
  const uint32_t sV = sK + 2 * kv_bytes;                   // [2 sets][nd atoms]
  const int item0 = blockIdx.x * p.ipc;
  const int item1 = min(p.items, item0 + p.ipc);
  const int nitems = max(0, item1 - item0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQ);
    tma_prefetch_desc(&p.tmK);
    tma_prefetch_desc(&p.tmV);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kXMaxPipes; ++i) {
      mbar_init(bar0 + 8 * (kXQFull + i), 1);
      mbar_init(bar0 + 8 * (kXQFree + i), 1);
      mbar_init(bar0 + 8 * (kXSFull + i), 1);
      mbar_init(bar0 + 8 * (kXPFull + i), 128);
      mbar_init(bar0 + 8 * (kXOFull + i), 1);
      mbar_init(bar0 + 8 * (kXPipeFree + i), 128);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar0 + 8 * (kXKVFull + i), 1);
      mbar_init(bar0 + 8 * (kXKVFree + i), 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(smem_u32(&tmem_base_s), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_s, 0);
  pdl_launch();  // set-up done (common.cuh: programmatic dependent launch)

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    pdl_wait();  // q comes from the previous kernel of the step
    const uint32_t leader = elect_one();
    int run = -1, prev_bh = -1;
    for (int k = 0; k < nitems; ++k) {
      const int item = item0 + k;
      const int bh = item / p.qtiles, qt = item - bh * p.qtiles;
      const int b = bh / p.heads, head = bh - b * p.heads;
      if (bh != prev_bh) {
        prev_bh = bh;
        ++run;
        const int set = run & 1;
        mbar_wait(bar0 + 8 * (kXKVFree + set), ((run >> 1) & 1) ^ 1u);  // the run that used this set two runs ago is done
        const uint32_t bar = bar0 + 8 * (kXKVFull + set);
        mbar_arrive_expect_tx_w(leader, bar, 2 * kv_bytes);
        for (int a = 0; a < p.nd; ++a) {
          tma_load_4d_w(leader, sK + set * kv_bytes + a * kv_atom, &p.tmK, bar, a * 64, head, 0, b);
          tma_load_4d_w(leader, sV + set * kv_bytes + a * kv_atom, &p.tmV, bar, a * 64, head, 0, b);
        }
      }
      const int pipe = k % p.np, use = k / p.np;
      mbar_wait(bar0 + 8 * (kXQFree + pipe), (use & 1) ^ 1u);
      const uint32_t qbar = bar0 + 8 * (kXQFull + pipe);
      mbar_arrive_expect_tx_w(leader, qbar, q_bytes);
      for (int a = 0; a < p.nd; ++a)
        tma_load_4d_w(leader, sQ + pipe * q_bytes + a * kXQAtom, &p.tmQ, qbar, a * 64, head, qt * kXBQ, b);
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer: a polling scheduler over two
    // streams of work - S_k (needs Q_k, the run's K, pipe k % np free) and O_k (needs P_k) - so neither waits behind
    // the other.  S may run at most np items ahead of O.
    const uint32_t leader = elect_one();
    const uint32_t idesc_s = make_idesc_f16(kXBQ, p.KP, 0, 0);
    const uint32_t idesc_o = make_idesc_f16(kXBQ, p.dN, 0, 1);
    const uint64_t dQ = make_smem_desc_sw128(sQ, 16, 1024);
    const uint64_t dK = make_smem_desc_sw128(sK, 16, 1024);
    const uint64_t dV = make_smem_desc_sw128(sV, kv_atom, 1024);  // MN-major: d atoms kv_atom apart
    const uint64_t dP = make_smem_desc_sw128(sP, 16, 1024);
    const uint32_t q_lo = static_cast<uint32_t>(dQ), q_hi = static_cast<uint32_t>(dQ >> 32);
    const uint32_t k_lo = static_cast<uint32_t>(dK), k_hi = static_cast<uint32_t>(dK >> 32);
    const uint32_t v_lo = static_cast<uint32_t>(dV), v_hi = static_cast<uint32_t>(dV >> 32);
    const uint32_t p_lo = static_cast<uint32_t>(dP), p_hi = static_cast<uint32_t>(dP >> 32);
    const int ksteps_s = p.dN / 16, ksteps_o = p.KP / 16;
    // per-item run bookkeeping is recomputed on both streams (cheap integer work)
    int ks = 0, ko = 0;          // next S / next O item
    int run_s = -1, bh_s = -1;   // S stream
    int run_o = -1, bh_o = -1;   // O stream
    uint32_t idle = 0;
    long long t0 = 0;
    while (ko < nitems) {
      bool progress = false;
      if (ks < nitems && ks < ko + p.np) {
        const int item = item0 + ks;
        const int bh = item / p.qtiles;
        const int run = (bh != bh_s) ? run_s + 1 : run_s;
        const int set = run & 1, pipe = ks % p.np, use = ks / p.np;
        if (mbar_test(bar0 + 8 * (kXQFull + pipe), use & 1) && mbar_test(bar0 + 8 * (kXKVFull + set), (run >> 1) & 1) &&
            mbar_test(bar0 + 8 * (kXPipeFree + pipe), (use & 1) ^ 1u)) {
          tc_fence_after();
          const uint32_t a0 = q_lo + pipe * (q_bytes >> 4), b0 = k_lo + set * (kv_bytes >> 4);
          const uint32_t tS = tmem_base + pipe * p.tstride;
          for (int s = 0; s < ksteps_s; ++s)
            umma_f16_ss_w(leader, tS, a0 + (s >> 2) * (kXQAtom >> 4) + (s & 3) * 2, q_hi,
                          b0 + (s >> 2) * (kv_atom >> 4) + (s & 3) * 2, k_hi, idesc_s, s != 0);
          umma_commit_w(leader, bar0 + 8 * (kXSFull + pipe));
          umma_commit_w(leader, bar0 + 8 * (kXQFree + pipe));
          run_s = run;
          bh_s = bh;
          ++ks;
          progress = true;
        }
      }
      if (ko < ks) {
        const int item = item0 + ko;
        const int bh = item / p.qtiles;
        const int run = (bh != bh_o) ? run_o + 1 : run_o;
        const int set = run & 1, pipe = ko % p.np, use = ko / p.np;
        if (mbar_test(bar0 + 8 * (kXPFull + pipe), use & 1)) {
          tc_fence_after();
          const uint32_t a0 = p_lo + pipe * p_atoms * (kXQAtom >> 4), b0 = v_lo + set * (kv_bytes >> 4);
          const uint32_t tO = tmem_base + pipe * p.tstride + p.KP;
          for (int s = 0; s < ksteps_o; ++s)  // 16 keys per step: +32 B along P's K, +16 rows (2048 B) down V
            umma_f16_ss_w(leader, tO, a0 + (s >> 2) * (kXQAtom >> 4) + (s & 3) * 2, p_hi, b0 + s * 128, v_hi, idesc_o,
                          s != 0);
          umma_commit_w(leader, bar0 + 8 * (kXOFull + pipe));
          // last item of its run: the K/V set may be overwritten once this product has completed
          const bool last_of_run = (ko + 1 == nitems) || ((item + 1) / p.qtiles != bh);
          if (last_of_run) umma_commit_w(leader, bar0 + 8 * (kXKVFree + set));
          run_o = run;
          bh_o = bh;
          ++ko;
          progress = true;
        }
      }
      if (progress) {
        idle = 0;
        t0 = 0;
      } else if (((++idle) & 4095u) == 0u) {  // bounded like every other wait in this library
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > B200SD_WAIT_TIMEOUT_CYCLES) mbar_timeout_trap(bar0, 999u);
      }
    }
  } else if (warp >= 4 && (warp - 4) / 4 < p.np) {
    // ---------------------------------------------------------------- softmax warpgroup of pipeline `pipe`
    pdl_wait();  // the output rows may still be read by an earlier kernel
    const int pipe = (warp - 4) >> 2;
    const int qd = warp & 3;
    const int r = qd * 32 + lane;  // query row in the tile == TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
    const uint32_t tS = tmem_base + pipe * p.tstride + lane_off;
    const uint32_t tO = tS + p.KP;
    const uint32_t sPr = sP + pipe * p_atoms * kXQAtom + r * 128;
    const uint32_t swz = static_cast<uint32_t>(r & 7);
    const int nchunks = p.KP / 16;  // 16-column chunks of S (at most 8)
    for (int k = pipe; k < nitems; k += p.np) {
      const int use = k / p.np;
      const int item = item0 + k;
      const int bh = item / p.qtiles, qt = item - bh * p.qtiles;
      const int b = bh / p.heads, head = bh - b * p.heads;
      const int qrow = qt * kXBQ + r;
      mbar_wait_spin(bar0 + 8 * (kXSFull + pipe), use & 1);
      tc_fence_after();
      // the whole score row: up to 128 fp32 (only the chunks that exist are loaded)
      uint32_t sc[kCh][16];
#pragma unroll
      for (int c = 0; c < kCh; ++c)
        if (c < nchunks) tmem_ld16(tS + c * 16, sc[c]);
      tmem_ld_wait();
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < kCh; ++c)
        if (c < nchunks) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (c * 16 + i >= p.nk) sc[c][i] = __float_as_uint(-INFINITY);  // padding keys
            mx = fmaxf(mx, __uint_as_float(sc[c][i]));
          }
        }
      const float mb = mx * p.scale_log2;
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < kCh; ++c)
        if (c < nchunks) {
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float p0 = x_ex2(fmaf(__uint_as_float(sc[c][2 * i]), p.scale_log2, -mb));
            const float p1 = x_ex2(fmaf(__uint_as_float(sc[c][2 * i + 1]), p.scale_log2, -mb));
            sum += p0 + p1;
            w[i] = pack_act2(p0, p1);
          }
          // columns [16c, 16c + 16) = 16-byte chunks 2c, 2c + 1 of my row: atom (2c) / 8, chunk (2c) % 8 (swizzled)
          const uint32_t at = sPr + static_cast<uint32_t>(c >> 2) * kXQAtom;
          const uint32_t ch = static_cast<uint32_t>((2 * c) & 7);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at + ((ch ^ swz) << 4)), "r"(w[0]), "r"(w[1]),
                       "r"(w[2]), "r"(w[3])
                       : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at + (((ch + 1) ^ swz) << 4)), "r"(w[4]),
                       "r"(w[5]), "r"(w[6]), "r"(w[7])
                       : "memory");
        }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar0 + 8 * (kXPFull + pipe));
      const float inv = 1.f / sum;
      mbar_wait_spin(bar0 + 8 * (kXOFull + pipe), use & 1);
      tc_fence_after();
      act_t* orow = p.out + (static_cast<size_t>(b) * p.nq + qrow) * p.ldo + head * p.d;
      for (int c = 0; c < p.dN; c += 16) {
        uint32_t o[16];
        tmem_ld16(tO + c, o);
        tmem_ld_wait();
        if (qrow < p.nq) {
#pragma unroll
          for (int h8 = 0; h8 < 2; ++h8) {
            if (c + h8 * 8 < p.d) {  // d is a multiple of 8
              uint4 u;
              u.x = pack_act2(__uint_as_float(o[h8 * 8 + 0]) * inv, __uint_as_float(o[h8 * 8 + 1]) * inv);
              u.y = pack_act2(__uint_as_float(o[h8 * 8 + 2]) * inv, __uint_as_float(o[h8 * 8 + 3]) * inv);
              u.z = pack_act2(__uint_as_float(o[h8 * 8 + 4]) * inv, __uint_as_float(o[h8 * 8 + 5]) * inv);
              u.w = pack_act2(__uint_as_float(o[h8 * 8 + 6]) * inv, __uint_as_float(o[h8 * 8 + 7]) * inv);
              *reinterpret_cast<uint4*>(orow + c + h8 * 8) = u;
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(bar0 + 8 * (kXPipeFree + pipe));  // S / O columns and the P tile may be reused
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

struct XAttnLaunch {
  XAttnDev d;
  int grid, threads, smem_bytes;
};

// nk <= 128, not causal: the persistent short-KV kernel.  Returns B200SD_OK after launching, or a positive value when
// the shape does not fit (the caller then uses the general kernel).
int launch_cross_attention(const b200sd_attn_params& h, const DeviceInfo& di, cudaStream_t stream) {
  if (int e = ensure_dyn_smem(reinterpret_cast<const void*>(&xattn_kernel<5, 3>), di.max_smem_optin - 512)) return e;
  if (int e = ensure_dyn_smem(reinterpret_cast<const void*>(&xattn_kernel<8, 2>), di.max_smem_optin - 512)) return e;
  auto launch = [&](const XAttnLaunch& X) -> int {
    if (X.d.KP <= 80)
      B200SD_CUDA_CHECK(launch_pdl(kPdlAttention, xattn_kernel<5, 3>, dim3(X.grid), dim3(X.threads), X.smem_bytes, stream, X.d));
    else
      B200SD_CUDA_CHECK(launch_pdl(kPdlAttention, xattn_kernel<8, 2>, dim3(X.grid), dim3(X.threads), X.smem_bytes, stream, X.d));
    return B200SD_OK;
  };
  XAttnLaunch L;
  if (launch_cache_get(2, &h, sizeof(h), &L, sizeof(L))) return launch(L);
  memset(&L, 0, sizeof(L));
  XAttnDev& d = L.d;
  d.out = static_cast<act_t*>(h.out);
  d.batch = h.batch;
  d.heads = h.heads;
  d.nq = h.nq;
  d.nk = h.nk;
  d.d = h.d;
  d.ldo = h.ldo;
  d.nd = ceil_div(h.d, 64);
  d.dN = ceil_div(h.d, 16) * 16;
  d.KP = ceil_div(h.nk, 16) * 16;
  d.tstride = ceil_div(d.KP + d.dN, 32) * 32;
  d.qtiles = ceil_div(h.nq, kXBQ);
  d.items = h.batch * h.heads * d.qtiles;
  d.scale_log2 = h.scale * 1.4426950408889634f;
  // pipelines: as many as TMEM (512 columns) and shared memory allow, at most kXMaxPipes
  const int kv_bytes = d.nd * d.KP * 128;
  const int per_pipe = d.nd * static_cast<int>(kXQAtom) + ceil_div(d.KP, 64) * static_cast<int>(kXQAtom);
  const int budget = di.max_smem_optin - 2048 - 4 * kv_bytes;
  d.np = min(d.KP <= 80 ? 3 : 2, min(512 / d.tstride, budget / per_pipe));
  if (d.np < 1) return 1;
  L.grid = min(d.items, di.sm_count);
  d.ipc = ceil_div(d.items, L.grid);
  L.grid = ceil_div(d.items, d.ipc);
  L.threads = 128 + 128 * d.np;
  L.smem_bytes = 4 * kv_bytes + d.np * per_pipe + 1024;

  auto enc = [&](CUtensorMap* tm, const void* ptr, int rows, int ld, uint32_t box_rows) -> int {
    const uint64_t dims[4] = {static_cast<uint64_t>(h.d), static_cast<uint64_t>(h.heads), static_cast<uint64_t>(rows),
                              static_cast<uint64_t>(h.batch)};
    const uint64_t st[3] = {static_cast<uint64_t>(h.d) * 2, static_cast<uint64_t>(ld) * 2,
                            static_cast<uint64_t>(ld) * 2 * rows};
    const uint32_t box[4] = {64, 1, box_rows, 1};
    return encode_tmap_act(tm, ptr, 4, dims, st, box);
  };
  if (int e = enc(&d.tmQ, h.q, h.nq, h.ldq, kXBQ)) return e;
  if (int e = enc(&d.tmK, h.k, h.nk, h.ldk, static_cast<uint32_t>(d.KP))) return e;
  if (int e = enc(&d.tmV, h.v, h.nk, h.ldv, static_cast<uint32_t>(d.KP))) return e;
  launch_cache_put(2, &h, sizeof(h), &L, sizeof(L));
  return launch(L);
}

}  // namespace b200sd
