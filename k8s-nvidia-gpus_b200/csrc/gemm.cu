// Implicit-GEMM convolution / linear layer on tcgen05 (sm_100a).
//
// One persistent, warp-specialised kernel serves every dense contraction of the hot path except attention:
// 3x3 and 1x1 convolutions over NHWC activations and plain linear layers.
//
//   warp 0 (1 thread) : TMA producer.  The A operand is never materialised as an im2col matrix: for filter tap
//                       (kh, kw) a 4-D TMA box {64 channels, bw, bh, bn} is fetched at pixel offset
//                       (w0 + kw - pad, h0 + kh - pad); out-of-bounds elements are zero-filled by the TMA unit,
//                       which implements the convolution's zero padding, ragged M tiles and ragged channel counts.
//                       Up to two NHWC sources are walked back to back along K (virtual channel concat).
//   warp 1 (1 thread) : tcgen05.mma issuer (128 x BN x 16 UMMAs, act_t -> fp32 accumulators in TMEM).
//   warp 2            : TMEM allocator (512 columns = 2 accumulator stages of up to 256 columns).
//   warps 4..7        : epilogue.  tcgen05.ld -> registers -> alpha/bias/time-embedding/activation/GEGLU/residual
//                       -> act_t or fp32 global stores.  Double-buffered accumulators let the epilogue of tile i
//                       overlap the mainloop of tile i+1.
//
// Shared memory: num_stages x (A 128x64 act_t = 16 KB | B BNx64 act_t) in the canonical K-major SWIZZLE_128B layout,
// filled by TMA and consumed directly by UMMA shared-memory descriptors.
#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kBlockM = 128;
static constexpr int kBlockK = 64;
static constexpr int kMaxStages = 8;
static constexpr int kGemmThreads = 256;
static constexpr uint32_t kABytes = kBlockM * kBlockK * 2;

struct alignas(64) GemmDev {
  CUtensorMap tmA0;
  CUtensorMap tmA1;
  CUtensorMap tmB;
  void* out;
  const float* bias;
  const float* rowbias;
  const act_t* residual;
  int M_total, N, BN;
  int nimg, H, W;
  int bw, bh, bn;
  int tiles_w, tiles_h, tiles_n, tiles_m, tiles_ncol;
  int ksize, pad, taps;
  int cblk0, cblk1;
  int num_stages;
  int ldc, ldr, ld_rowbias;
  int act, out_f32;
  float alpha;
};

__device__ __forceinline__ void decode_tile(const GemmDev& p, int tile, int& n0, int& h0, int& w0, int& col0) {
  const int mt = tile / p.tiles_ncol;
  const int nt = tile - mt * p.tiles_ncol;
  const int tw = mt % p.tiles_w;
  const int th = (mt / p.tiles_w) % p.tiles_h;
  const int tn = mt / (p.tiles_w * p.tiles_h);
  n0 = tn * p.bn;
  h0 = th * p.bh;
  w0 = tw * p.bw;
  col0 = nt * p.BN;
}

__global__ void __launch_bounds__(kGemmThreads, 1) gemm_conv_kernel(const __grid_constant__ GemmDev p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t tiles_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_bytes = static_cast<uint32_t>(p.BN) * kBlockK * 2;
  const uint32_t stage_bytes = kABytes + b_bytes;
  const int num_tiles = p.tiles_m * p.tiles_ncol;
  const int cblk = p.cblk0 + p.cblk1;
  const int num_kb = p.taps * cblk;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA0);
    if (p.cblk1 > 0) tma_prefetch_desc(&p.tmA1);
    tma_prefetch_desc(&p.tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&tfull_bar[s]), 1);
      mbar_init(smem_u32(&tempty_bar[s]), 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(smem_u32(&tmem_base_s), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int n0, h0, w0, col0;
        decode_tile(p, tile, n0, h0, w0, col0);
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / cblk;
          const int cb = kb - tap * cblk;
          const int kh = tap / p.ksize;
          const int kw = tap - kh * p.ksize;
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          const uint32_t bar = smem_u32(&full_bar[stage]);
          const uint32_t sa = tiles_base + stage * stage_bytes;
          mbar_arrive_expect_tx(bar, stage_bytes);
          if (cb < p.cblk0)
            tma_load_4d(sa, &p.tmA0, bar, cb * kBlockK, w0 + kw - p.pad, h0 + kh - p.pad, n0);
          else
            tma_load_4d(sa, &p.tmA1, bar, (cb - p.cblk0) * kBlockK, w0 + kw - p.pad, h0 + kh - p.pad, n0);
          tma_load_2d(sa + kABytes, &p.tmB, bar, kb * kBlockK, col0);
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ------------------------------------------------------------ MMA issuer
      const uint32_t idesc = make_idesc_f16(kBlockM, p.BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t sa = tiles_base + stage * stage_bytes;
          const uint64_t adesc = make_smem_desc_sw128(sa, 16, 1024);
          const uint64_t bdesc = make_smem_desc_sw128(sa + kABytes, 16, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // advancing K by 16 act_t = 32 bytes inside the 128-byte swizzle row: +2 in the (addr >> 4) field
            umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(smem_u32(&empty_bar[stage]));
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(smem_u32(&tfull_bar[acc]));
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue (128 threads, thread = accumulator row)
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int iw = r % p.bw;
    const int ih = (r / p.bw) % p.bh;
    const int in = r / (p.bw * p.bh);
    const bool geglu = (p.act == B200SD_ACT_GEGLU);
    const int n_out = geglu ? p.N / 2 : p.N;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int n0, h0, w0, col0;
      decode_tile(p, tile, n0, h0, w0, col0);
      const int img = n0 + in, y = h0 + ih, x = w0 + iw;
      const bool row_ok = (img < p.nimg) && (y < p.H) && (x < p.W);
      const long long row = (static_cast<long long>(img) * p.H + y) * p.W + x;
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * 256 + (static_cast<uint32_t>(q * 32) << 16);
      const int ncols = min(p.BN, p.N - col0);  // valid accumulator columns in this tile
      for (int c = 0; c < ncols; c += (geglu ? 64 : 32)) {
        uint32_t v[32];
        float f[32];
        tmem_ld32(taddr + c, v);
        tmem_ld_wait();
        const int col = col0 + c;
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
        if (p.bias != nullptr) {
          if (col + 32 <= p.N) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col) + j);
              f[4 * j + 0] += b4.x;
              f[4 * j + 1] += b4.y;
              f[4 * j + 2] += b4.z;
              f[4 * j + 3] += b4.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col + j < p.N) f[j] += __ldg(p.bias + col + j);
          }
        }
        int ocol = col;  // output column of f[0]
        if (geglu) {
          // columns [c, c+32) = value, [c+32, c+64) = gate (weights packed that way): out = value * gelu(gate)
          uint32_t g[32];
          tmem_ld32(taddr + c + 32, g);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float gv = __uint_as_float(g[j]) * p.alpha;
            if (p.bias != nullptr) gv += __ldg(p.bias + col + 32 + j);
            f[j] *= gelu_erf_f(gv);
          }
          ocol = col / 2;
        } else {
          if (p.rowbias != nullptr && row_ok) {
            const float* rb = p.rowbias + static_cast<long long>(img) * p.ld_rowbias + col;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col + j < p.N) f[j] += __ldg(rb + j);
          }
          if (p.act == B200SD_ACT_SILU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = silu_f(f[j]);
          } else if (p.act == B200SD_ACT_QUICK_GELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = quick_gelu_f(f[j]);
          }
        }
        if (row_ok) {
          const bool full = (ocol + 32 <= n_out);
          if (p.residual != nullptr) {
            const act_t* rp = p.residual + row * p.ldr + ocol;
            if (full && (p.ldr & 7) == 0) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(rp) + j);
                const float2 a = unpack_act2(u.x), b = unpack_act2(u.y), cc = unpack_act2(u.z),
                             d = unpack_act2(u.w);
                f[8 * j + 0] += a.x;
                f[8 * j + 1] += a.y;
                f[8 * j + 2] += b.x;
                f[8 * j + 3] += b.y;
                f[8 * j + 4] += cc.x;
                f[8 * j + 5] += cc.y;
                f[8 * j + 6] += d.x;
                f[8 * j + 7] += d.y;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (ocol + j < n_out) f[j] += from_act(rp[j]);
            }
          }
          if (p.out_f32) {
            float* op = reinterpret_cast<float*>(p.out) + row * p.ldc + ocol;
            if (full && (p.ldc & 3) == 0) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                reinterpret_cast<float4*>(op)[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (ocol + j < n_out) op[j] = f[j];
            }
          } else {
            act_t* op = reinterpret_cast<act_t*>(p.out) + row * p.ldc + ocol;
            if (full && (p.ldc & 7) == 0) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint4 u;
                u.x = pack_act2(f[8 * j + 0], f[8 * j + 1]);
                u.y = pack_act2(f[8 * j + 2], f[8 * j + 3]);
                u.z = pack_act2(f[8 * j + 4], f[8 * j + 5]);
                u.w = pack_act2(f[8 * j + 6], f[8 * j + 7]);
                reinterpret_cast<uint4*>(op)[j] = u;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (ocol + j < n_out) op[j] = to_act(f[j]);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1u;
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

// ---------------------------------------------------------------------------------------------------- host
static int pick_bn(int M_tiles, int N, int sm_count, bool geglu) {
  // Candidates are multiples of 32 (64 for GEGLU pairs).  Score = padded work / SM-wave efficiency.
  const int cands[] = {256, 192, 160, 128, 96, 64, 32};
  int best = 128;
  double best_cost = 1e30;
  for (int bn : cands) {
    if (geglu && (bn % 64) != 0) continue;
    if (bn > 32 && bn - 32 >= N) continue;  // tile far wider than N
    const int tn = (N + bn - 1) / bn;
    const long long tiles = static_cast<long long>(M_tiles) * tn;
    const long long waves = (tiles + sm_count - 1) / sm_count;
    // per-tile time ~ bn (MMA floor) with a fixed per-k-block overhead that favours wide tiles (smem bandwidth
    // bound below bn ~ 128 for single-CTA UMMA)
    const double tile_cost = static_cast<double>(bn < 128 ? 128 * 0.75 + bn * 0.25 : bn) + 8.0;
    const double cost = static_cast<double>(waves) * tile_cost;
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

static bool g_gemm_attr_set = false;

}  // namespace b200sd

using namespace b200sd;

extern "C" int b200sd_gemm(const b200sd_gemm_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "gemm: params is NULL");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DeviceInfo di;
  if (int e = get_device_info(&di)) return e;
  const b200sd_gemm_params& h = *hp;
  B200SD_REQUIRE(h.a0 && h.w && h.out, B200SD_ERR_BAD_ARGUMENT, "gemm: a0/w/out must be non-NULL");
  B200SD_REQUIRE(h.ksize == 1 || h.ksize == 3, B200SD_ERR_BAD_SHAPE, "gemm: ksize must be 1 or 3 (got %d)", h.ksize);
  B200SD_REQUIRE(h.nimg > 0 && h.h > 0 && h.w_ > 0 && h.c0 > 0 && h.n > 0 && h.c1 >= 0, B200SD_ERR_BAD_SHAPE,
                 "gemm: non-positive dimension");
  B200SD_REQUIRE((h.c0 % 8) == 0 && (h.c1 % 8) == 0 && (h.ld0 % 8) == 0 && (h.c1 == 0 || (h.ld1 % 8) == 0),
                 B200SD_ERR_BAD_SHAPE, "gemm: channel counts / pixel strides must be multiples of 8");
  B200SD_REQUIRE(h.ld0 >= h.c0 && (h.c1 == 0 || h.ld1 >= h.c1), B200SD_ERR_BAD_SHAPE, "gemm: ld < channels");
  B200SD_REQUIRE((h.c1 == 0) == (h.a1 == nullptr), B200SD_ERR_BAD_ARGUMENT, "gemm: a1 and c1 must agree");
  B200SD_REQUIRE(aligned16(h.a0) && aligned16(h.a1) && aligned16(h.w) && aligned16(h.out) && aligned16(h.bias) &&
                     aligned16(h.residual),
                 B200SD_ERR_BAD_ALIGNMENT, "gemm: pointers must be 16-byte aligned");
  const bool geglu = (h.act == B200SD_ACT_GEGLU);
  B200SD_REQUIRE(!geglu || (h.n % 64) == 0, B200SD_ERR_BAD_SHAPE, "gemm: GEGLU needs n %% 64 == 0");
  B200SD_REQUIRE(!geglu || (h.rowbias == nullptr && h.residual == nullptr), B200SD_ERR_BAD_ARGUMENT,
                 "gemm: GEGLU epilogue takes no rowbias/residual");
  B200SD_REQUIRE(h.act >= 0 && h.act <= 3, B200SD_ERR_BAD_ARGUMENT, "gemm: bad act %d", h.act);
  const int n_out = geglu ? h.n / 2 : h.n;
  B200SD_REQUIRE(h.ldc >= n_out, B200SD_ERR_BAD_SHAPE, "gemm: ldc (%d) < output columns (%d)", h.ldc, n_out);
  const long long M_total = static_cast<long long>(h.nimg) * h.h * h.w_;
  B200SD_REQUIRE(M_total < (1ll << 31), B200SD_ERR_TOO_LARGE, "gemm: M too large");

  GemmDev d;
  memset(&d, 0, sizeof(d));
  d.bw = min(kBlockM, pow2_ceil(h.w_));
  d.bh = min(kBlockM / d.bw, pow2_ceil(h.h));
  d.bn = kBlockM / (d.bw * d.bh);
  d.tiles_w = ceil_div(h.w_, d.bw);
  d.tiles_h = ceil_div(h.h, d.bh);
  d.tiles_n = ceil_div(h.nimg, d.bn);
  d.tiles_m = d.tiles_w * d.tiles_h * d.tiles_n;
  int bn_tile = h.bn_hint;
  if (bn_tile == 0) bn_tile = pick_bn(d.tiles_m, h.n, di.sm_count, geglu);
  B200SD_REQUIRE(bn_tile >= 32 && bn_tile <= 256 && (bn_tile % 32) == 0 && (!geglu || (bn_tile % 64) == 0),
                 B200SD_ERR_BAD_ARGUMENT, "gemm: bad bn_hint %d", bn_tile);
  d.BN = bn_tile;
  d.tiles_ncol = ceil_div(h.n, d.BN);
  d.M_total = static_cast<int>(M_total);
  d.N = h.n;
  d.nimg = h.nimg;
  d.H = h.h;
  d.W = h.w_;
  d.ksize = h.ksize;
  d.pad = h.ksize / 2;
  d.taps = h.ksize * h.ksize;
  d.cblk0 = ceil_div(h.c0, kBlockK);
  d.cblk1 = ceil_div(h.c1, kBlockK);
  const uint32_t stage_bytes = kABytes + d.BN * kBlockK * 2;
  const int smem_budget = di.max_smem_optin - 1024 /*align slack*/ - 512 /*static*/;
  d.num_stages = min(kMaxStages, smem_budget / static_cast<int>(stage_bytes));
  B200SD_REQUIRE(d.num_stages >= 2, B200SD_ERR_CUDA, "gemm: not enough shared memory (%d B opt-in)", di.max_smem_optin);
  d.out = h.out;
  d.bias = h.bias;
  d.rowbias = h.rowbias;
  d.residual = static_cast<const act_t*>(h.residual);
  d.ldc = h.ldc;
  d.ldr = h.ldr;
  d.ld_rowbias = h.ld_rowbias;
  d.act = h.act;
  d.out_f32 = (h.out_dtype == B200SD_F32);
  d.alpha = h.alpha;

  {
    const uint64_t dims[4] = {static_cast<uint64_t>(h.c0), static_cast<uint64_t>(h.w_), static_cast<uint64_t>(h.h),
                              static_cast<uint64_t>(h.nimg)};
    const uint64_t st[3] = {static_cast<uint64_t>(h.ld0) * 2, static_cast<uint64_t>(h.ld0) * 2 * h.w_,
                            static_cast<uint64_t>(h.ld0) * 2 * h.w_ * h.h};
    const uint32_t box[4] = {kBlockK, static_cast<uint32_t>(d.bw), static_cast<uint32_t>(d.bh),
                             static_cast<uint32_t>(d.bn)};
    if (int e = encode_tmap_act(&d.tmA0, h.a0, 4, dims, st, box)) return e;
  }
  if (h.c1 > 0) {
    const uint64_t dims[4] = {static_cast<uint64_t>(h.c1), static_cast<uint64_t>(h.w_), static_cast<uint64_t>(h.h),
                              static_cast<uint64_t>(h.nimg)};
    const uint64_t st[3] = {static_cast<uint64_t>(h.ld1) * 2, static_cast<uint64_t>(h.ld1) * 2 * h.w_,
                            static_cast<uint64_t>(h.ld1) * 2 * h.w_ * h.h};
    const uint32_t box[4] = {kBlockK, static_cast<uint32_t>(d.bw), static_cast<uint32_t>(d.bh),
                             static_cast<uint32_t>(d.bn)};
    if (int e = encode_tmap_act(&d.tmA1, h.a1, 4, dims, st, box)) return e;
  }
  {
    const uint64_t ktot = static_cast<uint64_t>(d.taps) * (d.cblk0 + d.cblk1) * kBlockK;
    const uint64_t dims[2] = {ktot, static_cast<uint64_t>(h.n)};
    const uint64_t st[1] = {ktot * 2};
    const uint32_t box[2] = {kBlockK, static_cast<uint32_t>(d.BN)};
    if (int e = encode_tmap_act(&d.tmB, h.w, 2, dims, st, box)) return e;
  }

  const int smem_bytes = d.num_stages * stage_bytes + 1024;
  if (!g_gemm_attr_set) {
    B200SD_CUDA_CHECK(
        cudaFuncSetAttribute(gemm_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, di.max_smem_optin - 512));
    g_gemm_attr_set = true;
  }
  const int num_tiles = d.tiles_m * d.tiles_ncol;
  const int grid = min(num_tiles, di.sm_count);
  gemm_conv_kernel<<<grid, kGemmThreads, smem_bytes, stream>>>(d);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}
