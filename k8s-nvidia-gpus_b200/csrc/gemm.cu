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
//   warp 1 (1 thread) : tcgen05.mma issuer (128 x N x 16 UMMAs, 16-bit operands -> fp32 accumulators in TMEM; N is
//                       the tile width, shrunk to the valid width on the last column tile).
//   warp 2            : TMEM allocator (512 columns = 2 accumulator stages of up to 256 columns).
//   warps 4..11       : epilogue, 2 warpgroups on alternate 128-byte column groups (64 16-bit / 32 fp32 columns).
//                       tcgen05.ld -> registers -> alpha/bias/time-embedding/activation/GEGLU/residual -> one
//                       swizzle-128B staging tile per warpgroup -> ONE 4-D TMA store per group (the TMA unit clips
//                       ragged rows/columns, so the epilogue has no address arithmetic or predicates).  The residual
//                       tile arrives the same way (TMA load into the staging tile).  Double-buffered accumulators
//                       let the epilogue of tile i overlap the mainloop of tile i+1.
//
// Shared memory: num_stages x (A 128x64 = 16 KB | B BNx64) in the canonical K-major SWIZZLE_128B layout, filled by
// TMA and consumed directly by UMMA shared-memory descriptors, + 2 x 16 KB epilogue staging.
#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kBlockM = 128;
static constexpr int kBlockK = 64;
static constexpr int kMaxStages = 8;
static constexpr int kGemmThreads = 384;
static constexpr uint32_t kABytes = kBlockM * kBlockK * 2;
static constexpr uint32_t kEpiTileBytes = 128 * 128;        // 128 rows x 128 B
static constexpr int kEpiStageBytes = 2 * kEpiTileBytes;    // one staging tile per epilogue warpgroup

// Division by a launch-time constant as multiply-high + shift (exact for dividends < 2^31): the tile decode runs once
// per tile in the epilogue's critical path, where four or five 32-bit integer divisions cost several hundred cycles.
// mul = ceil(2^(31 + s) / d), s = ceil(log2 d); d == 1 is flagged by shift < 0.
struct FastDiv {
  uint32_t mul;
  int32_t shift;
  uint32_t d;
};
static inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  if (d <= 1) {
    f.mul = 0;
    f.shift = -1;
    return f;
  }
  int s = 0;
  while ((1u << s) < d) ++s;
  f.mul = static_cast<uint32_t>(((1ull << (31 + s)) + d - 1) / d);
  f.shift = s - 1;
  return f;
}
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv& f) {
  return f.shift < 0 ? n : (__umulhi(n, f.mul) >> f.shift);
}

struct alignas(64) GemmDev {
  CUtensorMap tmA0;
  CUtensorMap tmA1;
  CUtensorMap tmB;
  CUtensorMap tmO;  // output, box {64 16-bit | 32 fp32 columns, bw, bh, bn}
  CUtensorMap tmR;  // residual (16-bit), same box
  const float* bias;
  const float* rowbias;
  unsigned long long* gn_stats;  // fused GroupNorm statistics of the output (32-byte fixed-point slots), or nullptr
  int gn_groups, gn_cpg, gn_rows_per_img;
  int N, BN;
  int nimg, H, W;
  int bw, bh, bn;
  int tiles_w, tiles_h, tiles_n, tiles_m, tiles_ncol;
  FastDiv div_munits, div_tw, div_twth;  // tile -> column tile; M tile -> (tw, th, tn)
  // Tile order.  0: M fastest (all M tiles of column tile 0, then column tile 1, ...), CTAs striding over the list.
  // 1: column fastest AND every CTA takes a contiguous run of the list - the column tiles of one M tile run back to
  // back on the same CTA, so an A operand larger than L2 is read from HBM once instead of once per column tile
  // (65536 x 1280 -> 320: 168 MB of activations, two column tiles), and every CTA still sees the same mix of wide
  // and narrow column tiles.  The loops keep their strided index; split_tile() maps it to the run.
  int col_fast;
  FastDiv div_ncol, div_ctas;  // column tiles; CTAs (pairs) of the grid
  int run_q, run_rem;          // tiles / CTAs: quotient and remainder (CTA b owns q + (b < rem) consecutive tiles)
#ifdef B200SD_DIAG
  int diag;  // timing experiments only (results are wrong): 1 no epilogue work, 2 no B loads, 4 no A loads, 8 no output store, 16 epilogue math only
#endif
  int kw_taps, taps;     // filter taps along w; total taps (kh * kw)
  int off_h, off_w;      // input pixel of tap (0, 0) relative to output pixel * stride (-pad for a centred filter)
  int stride;            // 1, or 2: the A tensor map then walks every other input pixel (element strides)
  int cblk0, cblk1;
  int num_stages;
  int res_bufs;  // staging tiles per epilogue warpgroup: 2 = the residual of the NEXT group is prefetched
  // wide tiles (256 < BN <= 512, single-CTA only): ONE accumulator of BN TMEM columns instead of two of 256, two MMAs
  // and two B loads per step.  For problems of about one wave (M = 4096 x N = 1280 makes 160 tiles of 128 x 256 for
  // 148 SMs: two rounds) a 128 x 320 tile brings the count under the SM count; nothing is lost by not double-buffering
  // an accumulator that is used once.
  int wide;
  int pair;  // 1: CTA pairs (cta_group::2): M tile 256 across two CTAs, each CTA loads half of B
  int ld_rowbias;
  int act;
  float alpha;
  // LayerNorm folded into the GEMM on both sides (b200sd.h: ln_*):
  //   consumer: out = rstd_m * acc_mn - rstd_m * mean_m * ln_colsum[n] + bias[n], (mean, rstd) of A's row m from ln_in
  //   producer: sum / sum of squares of every OUTPUT row added to ln_out (the next LayerNorm's input statistics)
  const unsigned long long* ln_in;
  unsigned long long* ln_out;
  const float* ln_colsum;
  float ln_inv_k, ln_eps;
  // split-K: the reduction of every tile is cut into split_k slices that run as separate work units (on different
  // CTAs when there are CTAs to spare); each unit writes its raw fp32 accumulator tile to `ws`, the unit that arrives
  // last at the tile's counter adds the slices IN INDEX ORDER (so the result does not depend on who arrives when) and
  // runs the epilogue.  Counters are zero at rest (the finishing unit re-arms its tile).
  int split_k;
  int b_static;  // weights are not written by anything in flight: they may be fetched before griddepcontrol.wait
  FastDiv div_split;
  float* ws;   // [tiles][split_k] fp32 slices of 128 x BN, each stored [BN / 32 chunks][8 pieces][128 rows][4]
  int* cnt;    // [tiles]
};

// tile -> (M tile, column tile), M fastest: with a persistent grid-stride loop every CTA then sees the same mix of
// full-width and (narrower) last-column tiles instead of always the same column.  In pair mode `tile` counts PAIR
// tiles (two consecutive M tiles, one per CTA).
// tile -> (M unit, column tile) in the launch's tile order
__device__ __forceinline__ void split_tile(const GemmDev& p, int tile, int& mt, int& nt) {
  if (p.col_fast) {
    const int k = static_cast<int>(fdiv(static_cast<uint32_t>(tile), p.div_ctas));  // my k-th tile
    const int b = tile - k * static_cast<int>(p.div_ctas.d);                        // my CTA (pair)
    const int t = b * p.run_q + min(b, p.run_rem) + k;
    mt = static_cast<int>(fdiv(static_cast<uint32_t>(t), p.div_ncol));
    nt = t - mt * p.tiles_ncol;
  } else {
    nt = static_cast<int>(fdiv(static_cast<uint32_t>(tile), p.div_munits));
    mt = tile - nt * static_cast<int>(p.div_munits.d);
  }
}
__device__ __forceinline__ int tile_col0(const GemmDev& p, int tile) {
  int mt, nt;
  split_tile(p, tile, mt, nt);
  return nt * p.BN;
}
__device__ __forceinline__ void decode_tile(const GemmDev& p, int tile, int rank, int& n0, int& h0, int& w0, int& col0) {
  int mt, nt;
  split_tile(p, tile, mt, nt);
  if (p.pair) mt = 2 * mt + rank;  // an M tile past the end is all out of bounds: TMA zero-fills / clips it
  const int tn = static_cast<int>(fdiv(static_cast<uint32_t>(mt), p.div_twth));
  const int rem = mt - tn * static_cast<int>(p.div_twth.d);
  const int th = static_cast<int>(fdiv(static_cast<uint32_t>(rem), p.div_tw));
  const int tw = rem - th * p.tiles_w;
  n0 = tn * p.bn;
  h0 = th * p.bh;
  w0 = tw * p.bw;
  col0 = nt * p.BN;
}

// mbarrier indices inside the kernel's barrier block
enum : int {
  kBarFull = 0,                        // [stage]  A/B tiles landed
  kBarEmpty = kBarFull + kMaxStages,   // [stage]  MMAs that read the stage have completed
  kBarTFull = kBarEmpty + kMaxStages,  // [acc]    accumulator complete
  kBarTEmpty = kBarTFull + 2,          // [acc]    accumulator drained by the epilogue
  kBarRes = kBarTEmpty + 2,            // [epilogue warpgroup] residual tile landed
  kNumBars = kBarRes + 4              // [epilogue warpgroup][staging buffer]
};

// kSimple: the epilogue of a plain linear / 1x1 layer (bias, GEGLU or residual only) - no split-K reduction, no alpha, no
// row bias, no SiLU / GELU, no folded LayerNorm, no fused GroupNorm statistics.  Those features are uniform run-time
// branches otherwise; compiled out they free registers and instructions in the epilogue that bounds the
// short-reduction layers.
template <bool kOutF32, bool kGeglu, bool kRes, bool kPair, bool kSimple>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_conv_kernel(const __grid_constant__ GemmDev p) {
  extern __shared__ uint8_t smem_raw[];
  // all mbarriers in one block: one base address, every barrier a constant byte offset from it
  __shared__ __align__(8) uint64_t bars[kNumBars];
  __shared__ uint32_t tmem_base_s;
  const uint32_t bar0 = smem_u32(bars);

  const int warp = warp_idx_uniform();
  const int lane = threadIdx.x & 31;
  const uint32_t tiles_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;  // 0 = leader (issues the MMAs, owns the full barriers)
  const uint32_t b_bytes = static_cast<uint32_t>(kPair ? p.BN / 2 : p.BN) * kBlockK * 2;  // this CTA's share of B
  const uint32_t stage_bytes = kABytes + b_bytes;
  const int nacc = (!kPair && p.wide) ? 1 : 2;  // TMEM accumulators in rotation
  const int num_tiles = (kPair ? (p.tiles_m + 1) / 2 : p.tiles_m) * p.tiles_ncol;
  const int num_units = num_tiles * p.split_k;  // work units of the persistent loop: (tile, K slice)
  const int tile0 = kPair ? blockIdx.x / 2 : blockIdx.x;
  const int tile_step = kPair ? gridDim.x / 2 : gridDim.x;
  const int cblk = p.cblk0 + p.cblk1;
  const int num_kb = p.taps * cblk;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA0);
    if (p.cblk1 > 0) tma_prefetch_desc(&p.tmA1);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmO);
    if (kRes) tma_prefetch_desc(&p.tmR);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init((bar0 + 8 * (kBarFull + (s))), 1);
      mbar_init((bar0 + 8 * (kBarEmpty + (s))), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init((bar0 + 8 * (kBarTFull + (s))), 1);
      mbar_init((bar0 + 8 * (kBarTEmpty + (s))), kPair ? 16 : 8);  // one arrive per epilogue warp (of both CTAs)
      mbar_init((bar0 + 8 * (kBarRes + 2 * s)), 1);
      mbar_init((bar0 + 8 * (kBarRes + 2 * s + 1)), 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if (kPair)
      tmem_alloc_pair(smem_u32(&tmem_base_s), 512);
    else
      tmem_alloc(smem_u32(&tmem_base_s), 512);
  }
  tc_fence_before();
  if (kPair)
    cluster_sync_all();  // both CTAs' barriers are initialised before any remote arrive / multicast commit
  else
    __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_s, 0);  // warp-uniform for the issue loop
  pdl_launch();  // set-up done: the next kernel's blocks may take the SMs this grid leaves (see common.cuh)

  if (warp == 0) {
    // -------------------------------------------------------------- TMA producer: whole warp converged, one elected
    // lane issues (see the "_w" note in common.cuh); the k-block -> (tap, channel block) walk is incremental
    const uint32_t leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    // Programmatic dependent launch: this block may be resident while the previous kernel of the stream still runs.
    // Static weights do not depend on it: the B tiles of the first ring pass are requested (and their barriers armed
    // for the whole stage) before the wait, the A tiles follow after it.
    int pre = 0;
    if (p.b_static && tile0 < num_units) {
      const int tile = static_cast<int>(fdiv(static_cast<uint32_t>(tile0), p.div_split));
      const int split = tile0 - tile * p.split_k;
      const int kb_begin = split * num_kb / p.split_k, kb_end = (split + 1) * num_kb / p.split_k;
      int n0, h0, w0, col0;
      decode_tile(p, tile, rank, n0, h0, w0, col0);
      const int ntile = (min(p.BN, p.N - col0) + 15) & ~15;
      const int bcol = kPair ? col0 + static_cast<int>(rank) * (ntile / 2) : col0;
      pre = min(p.num_stages, kb_end - kb_begin);
      for (int s = 0; s < pre; ++s) {
        const uint32_t bar = bar0 + 8 * (kBarFull + s);
        const uint32_t sb = tiles_base + s * stage_bytes + kABytes;
        const int kcol = (kb_begin + s) * kBlockK;
        if (kPair) {
          if (rank == 0) mbar_arrive_expect_tx_w(leader, bar, 2 * stage_bytes);
          tma_load_2d_pair_w(leader, sb, &p.tmB, bar, kcol, bcol);
        } else {
          mbar_arrive_expect_tx_w(leader, bar, stage_bytes);
          tma_load_2d_w(leader, sb, &p.tmB, bar, kcol, bcol);
          if (p.wide) tma_load_2d_w(leader, sb + b_bytes / 2, &p.tmB, bar, kcol, bcol + p.BN / 2);
        }
      }
    }
    pdl_wait();  // activations come from earlier kernels of the step
    for (int unit = tile0; unit < num_units; unit += tile_step) {
      const int tile = static_cast<int>(fdiv(static_cast<uint32_t>(unit), p.div_split));
      const int split = unit - tile * p.split_k;
      const int kb_begin = split * num_kb / p.split_k, kb_end = (split + 1) * num_kb / p.split_k;
      int n0, h0, w0, col0;
      decode_tile(p, tile, rank, n0, h0, w0, col0);
      // pair mode: this CTA supplies rows [rank * n/2, (rank + 1) * n/2) of the (possibly narrowed) column tile
      const int ntile = (min(p.BN, p.N - col0) + 15) & ~15;
      const int bcol = kPair ? col0 + static_cast<int>(rank) * (ntile / 2) : col0;
      const int tap0 = kb_begin / cblk;
      int cb = kb_begin - tap0 * cblk, kh = tap0 / p.kw_taps;
      int kw = tap0 - kh * p.kw_taps;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        const bool b_issued = pre > 0;  // first ring pass of the first unit: armed and B requested before the wait
        pre -= b_issued ? 1 : 0;
        mbar_wait(bar0 + 8 * (kBarEmpty + stage), phase ^ 1u);
        const uint32_t bar = bar0 + 8 * (kBarFull + stage);
        const uint32_t sa = tiles_base + stage * stage_bytes;
        const bool first = cb < p.cblk0;
        const CUtensorMap* tmA = first ? &p.tmA0 : &p.tmA1;
        const int ch = (first ? cb : cb - p.cblk0) * kBlockK;
        if (kPair) {
          if (rank == 0 && !b_issued) mbar_arrive_expect_tx_w(leader, bar, 2 * stage_bytes);  // both CTAs' loads land on the leader's barrier
          tma_load_4d_pair_w(leader, sa, tmA, bar, ch, w0 * p.stride + kw + p.off_w, h0 * p.stride + kh + p.off_h, n0);
          if (!b_issued) tma_load_2d_pair_w(leader, sa + kABytes, &p.tmB, bar, kb * kBlockK, bcol);
        } else if (b_issued) {
          tma_load_4d_w(leader, sa, tmA, bar, ch, w0 * p.stride + kw + p.off_w, h0 * p.stride + kh + p.off_h, n0);
#ifdef B200SD_DIAG
        } else if (p.diag & 6) {
          mbar_arrive_expect_tx_w(leader, bar, (p.diag & 4) ? b_bytes : kABytes);
          if (!(p.diag & 4))
            tma_load_4d_w(leader, sa, tmA, bar, ch, w0 * p.stride + kw + p.off_w, h0 * p.stride + kh + p.off_h, n0);
          if (!(p.diag & 2)) tma_load_2d_w(leader, sa + kABytes, &p.tmB, bar, kb * kBlockK, bcol);
#endif
        } else {
          mbar_arrive_expect_tx_w(leader, bar, stage_bytes);
          tma_load_4d_w(leader, sa, tmA, bar, ch, w0 * p.stride + kw + p.off_w, h0 * p.stride + kh + p.off_h, n0);
          if (p.wide) {  // box = BN / 2 rows
            tma_load_2d_w(leader, sa + kABytes, &p.tmB, bar, kb * kBlockK, bcol);
            tma_load_2d_w(leader, sa + kABytes + b_bytes / 2, &p.tmB, bar, kb * kBlockK, bcol + p.BN / 2);
          } else {
            tma_load_2d_w(leader, sa + kABytes, &p.tmB, bar, kb * kBlockK, bcol);
          }
        }
        if (++cb == cblk) {
          cb = 0;
          if (++kw == p.kw_taps) {
            kw = 0;
            ++kh;
          }
        }
        if (++stage == p.num_stages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      // ------------------------------------------------------------ MMA issuer (pair mode: the leader CTA only),
      // whole warp converged: descriptors stay on the uniform datapath, only the low words move
      const uint32_t leader = elect_one();
      const uint64_t desc0 = make_smem_desc_sw128(tiles_base, 16, 1024);
      const uint32_t d_lo = static_cast<uint32_t>(desc0), d_hi = static_cast<uint32_t>(desc0 >> 32);
      const uint32_t stage_u = stage_bytes >> 4, b_off = kABytes >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int unit = tile0; unit < num_units; unit += tile_step) {
        const int tile = static_cast<int>(fdiv(static_cast<uint32_t>(unit), p.div_split));
        const int split = unit - tile * p.split_k;
        const int kb_begin = split * num_kb / p.split_k, kb_end = (split + 1) * num_kb / p.split_k;
        const int col0 = tile_col0(p, tile);
        const int ncols = min(p.BN, p.N - col0);
        // no MMA work on padding columns; pair mode: one M = 256 instruction spans both CTAs
        const int ncols16 = (ncols + 15) & ~15;
        const uint32_t idesc = make_idesc_f16(kPair ? 2 * kBlockM : kBlockM, min(ncols16, 256), 0, 0);
        // wide tiles: columns 256.. take a second MMA per step (B rows 256.. sit 32 KB further in the stage)
        const int ncols_hi = ncols16 - 256;
        const uint32_t idesc_hi = make_idesc_f16(kBlockM, max(ncols_hi, 16), 0, 0);
        mbar_wait(bar0 + 8 * (kBarTEmpty + acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(bar0 + 8 * (kBarFull + stage), phase);
          tc_fence_after();
          const uint32_t a_lo = d_lo + stage * stage_u;
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // advancing K by 16 elements = 32 bytes inside the 128-byte swizzle row: +2 in the (addr >> 4) field
            if (kPair)
              umma_f16_ss_pair_w(leader, d_tmem, a_lo + 2 * k, d_hi, a_lo + b_off + 2 * k, d_hi, idesc, ((kb - kb_begin) | k) != 0);
            else
              umma_f16_ss_w(leader, d_tmem, a_lo + 2 * k, d_hi, a_lo + b_off + 2 * k, d_hi, idesc, ((kb - kb_begin) | k) != 0);
            if (!kPair && ncols_hi > 0)
              umma_f16_ss_w(leader, d_tmem + 256, a_lo + 2 * k, d_hi, a_lo + b_off + (256 * 128 >> 4) + 2 * k, d_hi,
                            idesc_hi, ((kb - kb_begin) | k) != 0);
          }
          if (kPair)
            umma_commit_pair_w(leader, bar0 + 8 * (kBarEmpty + stage));  // frees the stage in both CTAs
          else
            umma_commit_w(leader, bar0 + 8 * (kBarEmpty + stage));
          if (++stage == p.num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (kPair)
          umma_commit_pair_w(leader, bar0 + 8 * (kBarTFull + acc));
        else
          umma_commit_w(leader, bar0 + 8 * (kBarTFull + acc));
        if (++acc == nacc) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
    }
  } else if (warp == 3) {
    // -------------------------------------------------------------- weight prefetch (static weights only): the block
    // that owns M unit 0 of a column tile asks L2 for the rest of that tile's weight slab while the previous kernel of
    // the stream is still running (at small batch the step streams 1.7 GB of weights through short kernels that
    // would otherwise each start cold).  No dependence on earlier kernels, so no griddepcontrol.wait.
    if (p.b_static && tile0 < num_units) {
      const int tile = static_cast<int>(fdiv(static_cast<uint32_t>(tile0), p.div_split));
      const int split = tile0 - tile * p.split_k;
      int mt, nt;
      split_tile(p, tile, mt, nt);
      if (mt == 0) {
        const uint32_t leader = elect_one();
        const int kb_begin = split * num_kb / p.split_k, kb_end = (split + 1) * num_kb / p.split_k;
        const int col0 = nt * p.BN;
        const int ntile = (min(p.BN, p.N - col0) + 15) & ~15;
        const int bcol = kPair ? col0 + static_cast<int>(rank) * (ntile / 2) : col0;
        for (int kb = kb_begin + p.num_stages; kb < kb_end; ++kb) {
          tma_prefetch_l2_2d_w(leader, &p.tmB, kb * kBlockK, bcol);
          if (!kPair && p.wide) tma_prefetch_l2_2d_w(leader, &p.tmB, kb * kBlockK, bcol + p.BN / 2);
        }
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue: 2 warpgroups x 4 warps
    pdl_wait();  // row biases, residuals, statistics slots, split-K slices and the output all belong to the step
    const int q = warp & 3;            // TMEM lane quarter
    const int wg = (warp - 4) >> 2;    // warpgroup: alternate column groups
    const int r = q * 32 + lane;       // accumulator row == staging row
    const int iw = r % p.bw;
    const int ih = (r / p.bw) % p.bh;
    const int in = r / (p.bw * p.bh);
    constexpr int gcols = kOutF32 ? 32 : 64;        // output columns per 128-byte group
    constexpr int acols = kGeglu ? 2 * gcols : gcols;  // accumulator columns per group
    constexpr int nsub = kOutF32 ? 1 : 2;           // 32-column sub-blocks per group
    // staging: p.res_bufs tiles per warpgroup.  With two, group k of this warpgroup uses tile k & 1 and the residual
    // of group k + 1 is in flight while group k is computed and stored.
    const uint32_t stg0 = tiles_base + p.num_stages * stage_bytes + wg * p.res_bufs * kEpiTileBytes;
    const bool prefetch = kRes && p.res_bufs == 2;
    uint32_t gk = 0;  // groups this warpgroup has processed
    const uint32_t sw = static_cast<uint32_t>(r & 7);
    const bool issuer = (q == 0 && lane == 0);
    const int bar_id = 1 + wg;
    const uint32_t resb0 = bar0 + 8 * (kBarRes + 2 * wg);
    // (tile, group) after (tile, g) in this warpgroup's order; tile >= num_tiles when there is none
    auto next_group = [&](int& tile, int& g) {
      g += 2;
      for (;;) {
        if (tile >= num_tiles) return;
        const int c0 = tile_col0(p, tile);
        const int ng = (min(p.BN, p.N - c0) + acols - 1) / acols;
        if (g < ng) return;
        tile += tile_step;
        g = wg;
      }
    };
    auto load_residual = [&](int tile, int g, uint32_t buf) {  // issuer only
      int n0, h0, w0, col0;
      decode_tile(p, tile, rank, n0, h0, w0, col0);
      const int oc = kGeglu ? (col0 + g * acols) / 2 : col0 + g * acols;
      mbar_arrive_expect_tx(resb0 + 8 * buf, kEpiTileBytes);
      tma_load_4d(stg0 + buf * kEpiTileBytes, &p.tmR, resb0 + 8 * buf, oc, w0, h0, n0);
    };
    if (prefetch && issuer) {
      int t0 = tile0, g0 = wg - 2;
      next_group(t0, g0);
      if (t0 < num_tiles) load_residual(t0, g0, 0);
    }
    int acc = 0;
    uint32_t acc_phase = 0;
    __shared__ int split_last_s;
    for (int unit = tile0; unit < num_units; unit += tile_step) {
      const int tile = static_cast<int>(fdiv(static_cast<uint32_t>(unit), p.div_split));
      const int split = unit - tile * p.split_k;
      int n0, h0, w0, col0;
      decode_tile(p, tile, rank, n0, h0, w0, col0);
      const int img = min(n0 + in, p.nimg - 1);  // rows beyond the tensor are clipped by the TMA store
      // fused GroupNorm statistics: image of my warp's row slab, and whether my row exists at all
      const bool gn_row_ok = (n0 + in < p.nimg) && (h0 + ih < p.H) && (w0 + iw < p.W);
      const int gn_img = p.gn_rows_per_img > 0 ? (w0 + q * 32) / p.gn_rows_per_img : img;
      mbar_wait((bar0 + 8 * (kBarTFull + (acc))), acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * 256 + (static_cast<uint32_t>(q * 32) << 16);
      const int ncols = min(p.BN, p.N - col0);  // valid accumulator columns in this tile
#ifdef B200SD_DIAG
      const int ngroups = (p.diag & 1) ? 0 : (ncols + acols - 1) / acols;
#else
      const int ngroups = (ncols + acols - 1) / acols;
#endif

      // ---- split-K: park this unit's raw accumulator tile, count it, and go on unless it was the last slice
      const float* ws_tile = nullptr;
      if (!kSimple && p.split_k > 1) {
        // slice layout [32-column chunk][16-byte piece k][row] : the 32 lanes of a warp (consecutive rows) write 512
        // contiguous bytes per store (row-major rows would make every lane's 16 bytes a request of its own: measured,
        // the reduction then costs more than the split saves)
        const size_t slice_u4 = static_cast<size_t>(kBlockM) * p.BN / 4;
        uint4* wsl = reinterpret_cast<uint4*>(p.ws) + (static_cast<size_t>(tile) * p.split_k + split) * slice_u4 + r;
        for (int c = wg * 32; c < ncols; c += 64) {
          uint32_t v[32];
          tmem_ld32(taddr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 8; ++k)
            __stcg(wsl + ((c >> 5) * 8 + k) * kBlockM, make_uint4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive((bar0 + 8 * (kBarTEmpty + (acc))));  // the accumulator is free again
        if (++acc == nacc) {
          acc = 0;
          acc_phase ^= 1u;
        }
        __threadfence();  // my slice is visible device-wide before the tile's counter moves
        named_bar_sync(3, 256);
        if (warp == 4 && lane == 0) {
          const int old = atomicAdd(p.cnt + tile, 1);
          split_last_s = (old == p.split_k - 1);
          if (old == p.split_k - 1) p.cnt[tile] = 0;  // every slice has arrived: re-arm the counter for the next launch
        }
        named_bar_sync(3, 256);
        const bool last = split_last_s != 0;
        named_bar_sync(3, 256);  // (split_last_s may be rewritten by the next unit only after everyone has read it)
        if (!last) continue;
        __threadfence();
        ws_tile = p.ws + (static_cast<size_t>(tile) * p.split_k * slice_u4 + r) * 4;
      }

      // global row index of my accumulator row (row-major over [nimg, H, W]); used by the folded LayerNorm
      const long long grow = (static_cast<long long>(n0 + in) * p.H + (h0 + ih)) * p.W + (w0 + iw);
      // folded LayerNorm, consumer side: my row's mean and 1/std from the producer's integer sums
      float ln_a = 1.f, ln_b = 0.f;
      if (!kSimple && p.ln_in != nullptr && gn_row_ok) {
        const unsigned long long* sl = p.ln_in + grow * kGnSlotWords64;
        const ulonglong2 a2 = __ldcg(reinterpret_cast<const ulonglong2*>(sl));
        const ulonglong2 b2 = __ldcg(reinterpret_cast<const ulonglong2*>(sl + 2));
        const unsigned long long sw2[2] = {a2.x, a2.y}, qw2[2] = {b2.x, b2.y};
        const double m = fixed_value(sw2) * static_cast<double>(p.ln_inv_k);
        const double var = fmax(fixed_value(qw2) * static_cast<double>(p.ln_inv_k) - m * m, 0.0);
        ln_a = rsqrtf(static_cast<float>(var) + p.ln_eps);
        ln_b = -ln_a * static_cast<float>(m);
      }

      // accumulator columns [ac, ac+32) -> f[] = alpha * acc + bias
      auto load_chunk = [&](int ac, float(&f)[32]) {
        uint32_t v[32];
        if (!kSimple && ws_tile != nullptr) {
          // the slices of this tile, added in index order (deterministic whatever the arrival order was)
#pragma unroll
          for (int k = 0; k < 32; ++k) v[k] = 0u;
          for (int sl = 0; sl < p.split_k; ++sl) {
            const uint4* src = reinterpret_cast<const uint4*>(ws_tile) + static_cast<size_t>(sl) * (kBlockM * p.BN / 4) +
                               (ac >> 5) * 8 * kBlockM;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const uint4 u = __ldcg(src + k * kBlockM);
              v[4 * k + 0] = __float_as_uint(__uint_as_float(v[4 * k + 0]) + __uint_as_float(u.x));
              v[4 * k + 1] = __float_as_uint(__uint_as_float(v[4 * k + 1]) + __uint_as_float(u.y));
              v[4 * k + 2] = __float_as_uint(__uint_as_float(v[4 * k + 2]) + __uint_as_float(u.z));
              v[4 * k + 3] = __float_as_uint(__uint_as_float(v[4 * k + 3]) + __uint_as_float(u.w));
            }
          }
        } else {
          tmem_ld32(taddr + ac, v);
          tmem_ld_wait();
        }
        if (!kSimple && p.alpha != 1.f) {
#pragma unroll
          for (int k = 0; k < 32; ++k) f[k] = __uint_as_float(v[k]) * p.alpha;
        } else {
#pragma unroll
          for (int k = 0; k < 32; ++k) f[k] = __uint_as_float(v[k]);
        }
        if (!kSimple && p.ln_in != nullptr) {
          // LN(x) W^T = rstd * (x (gamma W)^T) - rstd * mean * colsum(gamma W)   (+ W beta + b, folded into bias)
          const int col = col0 + ac;
          if (col + 32 <= p.N) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.ln_colsum + col) + k);
              f[4 * k + 0] = fmaf(f[4 * k + 0], ln_a, ln_b * s4.x);
              f[4 * k + 1] = fmaf(f[4 * k + 1], ln_a, ln_b * s4.y);
              f[4 * k + 2] = fmaf(f[4 * k + 2], ln_a, ln_b * s4.z);
              f[4 * k + 3] = fmaf(f[4 * k + 3], ln_a, ln_b * s4.w);
            }
          } else {
#pragma unroll
            for (int k = 0; k < 32; ++k) {
              const float sn = (col + k < p.N) ? __ldg(p.ln_colsum + col + k) : 0.f;
              f[k] = fmaf(f[k], ln_a, ln_b * sn);
            }
          }
        }
        if (p.bias != nullptr) {
          const int col = col0 + ac;
          if (col + 32 <= p.N) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col) + k);
              fadd2(f[4 * k + 0], f[4 * k + 1], b4.x, b4.y, f[4 * k + 0], f[4 * k + 1]);
              fadd2(f[4 * k + 2], f[4 * k + 3], b4.z, b4.w, f[4 * k + 2], f[4 * k + 3]);
            }
          } else {
#pragma unroll
            for (int k = 0; k < 32; ++k)
              if (col + k < p.N) f[k] += __ldg(p.bias + col + k);
          }
        }
      };

      for (int g = wg; g < ngroups; g += 2) {
        const int ac0 = g * acols;
        const int ocol0 = kGeglu ? (col0 + ac0) / 2 : col0 + ac0;  // first output column of this group
        const uint32_t buf = prefetch ? (gk & 1u) : 0u;
        const uint32_t stg = stg0 + buf * kEpiTileBytes;
        const uint32_t my_row = stg + r * 128;
        const uint32_t resb = resb0 + 8 * buf;
        const uint32_t res_phase = prefetch ? ((gk >> 1) & 1u) : (gk & 1u);
        if (kRes) {
          // the residual tile lands in a staging tile (after the previous store from it has finished reading it)
          if (issuer) {
            bulk_wait_read0();
            if (prefetch) {
              int tn = tile, gn = g;
              next_group(tn, gn);
              if (tn < num_tiles) load_residual(tn, gn, buf ^ 1u);
            } else {
              mbar_arrive_expect_tx(resb, kEpiTileBytes);
              tma_load_4d(stg, &p.tmR, resb, ocol0, w0, h0, n0);
            }
          }
        }
        uint32_t w[32];  // this row's 128 output bytes
#pragma unroll
        for (int hh = 0; hh < nsub; ++hh) {
          float f[32];
          if (kGeglu) {
            // packed weights: 64-row groups [32 value | 32 gate] -> out = value * gelu(gate)
            float gt[32];
            load_chunk(ac0 + hh * 64, f);
            load_chunk(ac0 + hh * 64 + 32, gt);
#pragma unroll
            for (int k = 0; k < 32; k += 2) mul_gelu_erf2(f[k], f[k + 1], gt[k], gt[k + 1]);
          } else {
            const int ac = ac0 + hh * 32;
            load_chunk(ac, f);
            if (!kSimple && p.rowbias != nullptr) {
              const float* rb = p.rowbias + static_cast<long long>(img) * p.ld_rowbias + col0 + ac;
              if (col0 + ac + 32 <= p.N && (reinterpret_cast<uintptr_t>(rb) & 15u) == 0) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                  const float4 b4 = __ldg(reinterpret_cast<const float4*>(rb) + k);
                  fadd2(f[4 * k + 0], f[4 * k + 1], b4.x, b4.y, f[4 * k + 0], f[4 * k + 1]);
                  fadd2(f[4 * k + 2], f[4 * k + 3], b4.z, b4.w, f[4 * k + 2], f[4 * k + 3]);
                }
              } else {
#pragma unroll
                for (int k = 0; k < 32; ++k)
                  if (col0 + ac + k < p.N) f[k] += __ldg(rb + k);
              }
            }
            if (kSimple) {
            } else if (p.act == B200SD_ACT_SILU) {
#pragma unroll
              for (int k = 0; k < 32; ++k) f[k] = silu_f(f[k]);
            } else if (p.act == B200SD_ACT_QUICK_GELU) {
#pragma unroll
              for (int k = 0; k < 32; ++k) f[k] = quick_gelu_f(f[k]);
            } else if (p.act == B200SD_ACT_GELU) {
#pragma unroll
              for (int k = 0; k < 32; ++k) f[k] = gelu_erf_f(f[k]);
            }
          }
          if (kRes) {
            if (hh == 0) mbar_wait(resb, res_phase);
#pragma unroll
            for (int k = 0; k < 4; ++k) {  // my row, 16-byte chunks hh*4 .. hh*4+3
              uint4 u;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
                           : "r"(my_row + (((hh * 4 + k) ^ sw) << 4)));
              const float2 a = unpack_act2(u.x), b = unpack_act2(u.y), cc = unpack_act2(u.z), d = unpack_act2(u.w);
              fadd2(f[8 * k + 0], f[8 * k + 1], a.x, a.y, f[8 * k + 0], f[8 * k + 1]);
              fadd2(f[8 * k + 2], f[8 * k + 3], b.x, b.y, f[8 * k + 2], f[8 * k + 3]);
              fadd2(f[8 * k + 4], f[8 * k + 5], cc.x, cc.y, f[8 * k + 4], f[8 * k + 5]);
              fadd2(f[8 * k + 6], f[8 * k + 7], d.x, d.y, f[8 * k + 6], f[8 * k + 7]);
            }
          }
          if (kOutF32) {
#pragma unroll
            for (int k = 0; k < 32; ++k) w[k] = __float_as_uint(f[k]);
          } else {
#pragma unroll
            for (int k = 0; k < 16; ++k) w[hh * 16 + k] = pack_act2(f[2 * k], f[2 * k + 1]);
          }
        }
        if (!kOutF32) {
          if (!kSimple && p.ln_out != nullptr) {
            // folded LayerNorm, producer side: sum / sum of squares of my row's 64 (rounded) outputs of this group,
            // added as integers to the row's slot (order-independent across column groups and tiles)
            const int n_valid = (kGeglu ? p.N / 2 : p.N) - ocol0;
            float rs = 0.f, rq = 0.f;
#pragma unroll
            for (int k = 0; k < 32; ++k) {
              const float2 v = unpack_act2(w[k]);
              if (2 * k < n_valid) {
                rs += v.x + v.y;
                rq = fmaf(v.x, v.x, fmaf(v.y, v.y, rq));
              }
            }
            if (gn_row_ok) {
              unsigned long long* sl = p.ln_out + grow * kGnSlotWords64;
              fixed_add_global(sl, rs);
              fixed_add_global(sl + 2, rq);
            }
          }
        }
#ifdef B200SD_DIAG
        if (p.diag & 16) {  // compute only: no staging, no store (keep the packed words alive)
          if (w[0] == 0x7fc07fc0u && w[31] == 0x12345678u) p.cnt[0] = 1;
          ++gk;
          continue;
        }
#endif
        // staging tile free: residual consumed by everyone (kRes) / previous store has finished reading it
        if (!kRes && issuer) bulk_wait_read0();
        named_bar_sync(bar_id, 128);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((k ^ sw) << 4)), "r"(w[4 * k]),
                       "r"(w[4 * k + 1]), "r"(w[4 * k + 2]), "r"(w[4 * k + 3])
                       : "memory");
        fence_proxy_async_smem();
        if (!kOutF32 && !kGeglu) {
          if (!kSimple && p.gn_stats != nullptr) {
            // Fused GroupNorm statistics of this 128 x 64 output group, read back TRANSPOSED from the staging tile:
            // lane = one column pair, walking the 32 rows my own warp has just written (one image, see b200sd.h) -
            // conflict-free 4-byte loads, a rolled loop.  Lanes of the same channel group are then combined by a
            // segmented shuffle reduction and the segment heads add their sums to the (image, group) slot.
            __syncwarp();
            const uint32_t rows_ok = __ballot_sync(0xffffffffu, gn_row_ok);
            const uint32_t wofs = (static_cast<uint32_t>(lane) & 3u) * 4u, chunk = static_cast<uint32_t>(lane) >> 2;
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
              const uint32_t row = static_cast<uint32_t>(q * 32 + rr);
              uint32_t wv;
              asm volatile("ld.shared.b32 %0, [%1];" : "=r"(wv) : "r"(stg + row * 128 + ((chunk ^ (row & 7u)) << 4) + wofs));
              const float2 v = unpack_act2(wv);
              if ((rows_ok >> rr) & 1u) {
                s0 += v.x;
                q0 = fmaf(v.x, v.x, q0);
                s1 += v.y;
                q1 = fmaf(v.y, v.y, q1);
              }
            }
            const int col = ocol0 + 2 * lane;
            const int g = col / p.gn_cpg;  // channels per group are even: both columns of the pair share a group
            float gs = s0 + s1, gq = q0 + q1;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const float so = __shfl_down_sync(0xffffffffu, gs, o), qo = __shfl_down_sync(0xffffffffu, gq, o);
              const int go = __shfl_down_sync(0xffffffffu, g, o);
              if (lane + o < 32 && go == g) {
                gs += so;
                gq += qo;
              }
            }
            const int gprev = __shfl_up_sync(0xffffffffu, g, 1);
            if ((lane == 0 || gprev != g) && col < p.N) {
              // integer (fixed-point) accumulation: the total does not depend on which warp / CTA arrives first
              unsigned long long* gst = p.gn_stats + (static_cast<size_t>(gn_img) * p.gn_groups + g) * kGnSlotWords64;
              fixed_add_global(gst, gs);
              fixed_add_global(gst + 2, gq);
            }
          }
        }
        named_bar_sync(bar_id, 128);
#ifdef B200SD_DIAG
        if (issuer && !(p.diag & 8)) {
#else
        if (issuer) {
#endif
          tma_store_4d(&p.tmO, stg, ocol0, w0, h0, n0);
          bulk_commit();
        }
        ++gk;
      }
      if (!kSimple && p.split_k > 1) continue;  // (the accumulator was released right after its slice was parked)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kPair && rank != 0)
          mbar_arrive_remote((bar0 + 8 * (kBarTEmpty + (acc))), 0);  // the leader's MMA thread waits for both CTAs
        else
          mbar_arrive((bar0 + 8 * (kBarTEmpty + (acc))));
      }
      if (++acc == nacc) {
        acc = 0;
        acc_phase ^= 1u;
      }
    }
    if (issuer) bulk_wait0();  // all stores complete before the CTA (and its shared memory) goes away
  }

  tc_fence_before();
  if (kPair)
    cluster_sync_all();  // neither CTA may retire (or free TMEM) while its peer can still signal / read it
  else
    __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (kPair)
      tmem_dealloc_pair(tmem_base, 512);
    else
      tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------------------------- host
struct TilePlan {
  int bn;
  int pair;
  double cost;
};

// Tile widths are multiples of the epilogue group (`gran` accumulator columns).  Per 64-deep K block a CTA spends
// max(MMA time, operand fetch time): 128 x w x 64 MACs at 4096 MAC/cycle = 2w cycles for valid width w, against
// (128 + rows of B it loads) x 128 B of L2 -> SM traffic at the ~45 B/cycle/SM measured on this part = 2.84 cycles per
// row.  The fetch term dominates, hence wide tiles, and hence CTA pairs (cta_group::2: M = 256 across two CTAs, each
// loading only half of B) whenever there are enough M tiles to keep the pairs busy.
static TilePlan pick_tiles(int M_tiles, int N, int sm_count, int gran, int num_kb) {
  TilePlan best{gran, 0, 1e30};
  for (int pair = 0; pair < 2; ++pair) {
    if (pair && M_tiles < 2) continue;
    // (wide tiles - 256 < bn <= 512, bn_hint only - are not considered: measured on the one-wave shapes they were
    // meant for, M = 4096 x N = 1280, a 128 x 320 tile is within 6 % of two rounds of 128 x 256 either way, because
    // the single round then runs into the L2 -> SM operand bandwidth)
    for (int bn = 256; bn >= gran; bn -= gran) {
      if (bn - gran >= N && bn > gran) continue;  // wider than needed
      if (pair && (bn % 32) != 0) continue;       // each CTA's half of B must be whole 8-row swizzle groups (and /16)
      const int tn = (N + bn - 1) / bn;
      const long long units = static_cast<long long>(pair ? (M_tiles + 1) / 2 : M_tiles) * tn;
      const int slots = pair ? sm_count / 2 : sm_count;
      const long long waves = (units + slots - 1) / slots;
      const double w = static_cast<double>(bn < N ? bn : N);
      const double fetch = 2.84 * (128.0 + (pair ? bn / 2 : bn));
      const double per_kb = (2.0 * w > fetch ? 2.0 * w : fetch) + (pair ? 4000.0 : 600.0) / num_kb;  // pairs pay extra per-tile hand-offs
      const double cost = static_cast<double>(waves) * per_kb;
      if (cost < best.cost - 1e-9) best = TilePlan{bn, pair, cost};
    }
  }
  return best;
}

template <bool F, bool G, bool R, bool P, bool S>
static int launch_one(const GemmDev& d, int grid, int smem_bytes, int max_smem, cudaStream_t stream) {
  if (int e = ensure_dyn_smem(reinterpret_cast<const void*>(&gemm_conv_kernel<F, G, R, P, S>), max_smem - 512)) return e;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = P ? 2 : 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (pdl_enabled(kPdlGemm)) {  // programmatic dependent launch: set-up overlaps the previous kernel's tail
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.numAttrs = 2;
  }
  B200SD_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemm_conv_kernel<F, G, R, P, S>, d));
  return B200SD_OK;
}

template <bool F, bool G, bool R>
static int launch_variant(const GemmDev& d, int grid, int smem_bytes, int max_smem, cudaStream_t stream, bool simple) {
  if (!F && simple)
    return d.pair ? launch_one<false, G, R, true, true>(d, grid, smem_bytes, max_smem, stream)
                  : launch_one<false, G, R, false, true>(d, grid, smem_bytes, max_smem, stream);
  return d.pair ? launch_one<F, G, R, true, false>(d, grid, smem_bytes, max_smem, stream)
                : launch_one<F, G, R, false, false>(d, grid, smem_bytes, max_smem, stream);
}

// what a finished launch needs: the device-side descriptor block + launch geometry (kept in the launch cache)
struct GemmLaunch {
  GemmDev d;
  int grid, smem_bytes, variant;  // variant: 0 plain, 1 residual, 2 fp32 out, 3 GEGLU
  int simple;                     // the kSimple instantiation (see the kernel)
};

static int launch_cached(const GemmLaunch& L, int max_smem, cudaStream_t stream) {
  switch (L.variant) {
    case 3: return launch_variant<false, true, false>(L.d, L.grid, L.smem_bytes, max_smem, stream, L.simple != 0);
    case 2: return launch_variant<true, false, false>(L.d, L.grid, L.smem_bytes, max_smem, stream, L.simple != 0);
    case 1: return launch_variant<false, false, true>(L.d, L.grid, L.smem_bytes, max_smem, stream, L.simple != 0);
    default: return launch_variant<false, false, false>(L.d, L.grid, L.smem_bytes, max_smem, stream, L.simple != 0);
  }
}

}  // namespace b200sd

using namespace b200sd;

// Validation + tile planning + (unless ws_query is set) tensor maps and launch.  ws_query: only report the split-K
// workspace size for these parameters.
static int gemm_impl(const b200sd_gemm_params* hp, b200sd_stream stream_, size_t* ws_query) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "gemm: params is NULL");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DeviceInfo di;
  if (int e = get_device_info(&di)) return e;
  const b200sd_gemm_params& h = *hp;
  if (ws_query == nullptr) {
    // an identical parameter block was validated, planned and encoded before: replay it
    GemmLaunch hit;
    if (launch_cache_get(0, hp, sizeof(*hp), &hit, sizeof(hit))) return launch_cached(hit, di.max_smem_optin, stream);
  }
  B200SD_REQUIRE(h.a0 && h.w && h.out, B200SD_ERR_BAD_ARGUMENT, "gemm: a0/w/out must be non-NULL");
  B200SD_REQUIRE(h.ksize == 1 || h.ksize == 3, B200SD_ERR_BAD_SHAPE, "gemm: ksize must be 1 or 3 (got %d)", h.ksize);
  // filter geometry: ksize x ksize centred (default) or an explicit kh x kw window whose tap (0, 0) sits at input
  // pixel (out * stride + off_h, out * stride + off_w)
  const bool custom = (h.kh != 0 || h.kw != 0);
  const int f_kh = custom ? h.kh : h.ksize, f_kw = custom ? h.kw : h.ksize;
  const int f_offh = custom ? h.off_h : -(h.ksize / 2), f_offw = custom ? h.off_w : -(h.ksize / 2);
  const int f_stride = h.stride == 0 ? 1 : h.stride;
  B200SD_REQUIRE(f_kh >= 1 && f_kh <= 3 && f_kw >= 1 && f_kw <= 3, B200SD_ERR_BAD_SHAPE,
                 "gemm: filter window %d x %d unsupported (1..3)", f_kh, f_kw);
  B200SD_REQUIRE(f_stride == 1 || f_stride == 2, B200SD_ERR_BAD_SHAPE, "gemm: stride must be 1 or 2 (got %d)", h.stride);
  B200SD_REQUIRE(f_offh >= -3 && f_offh <= 3 && f_offw >= -3 && f_offw <= 3, B200SD_ERR_BAD_SHAPE, "gemm: bad tap origin");
  B200SD_REQUIRE(h.nimg > 0 && h.h > 0 && h.w_ > 0 && h.c0 > 0 && h.n > 0 && h.c1 >= 0, B200SD_ERR_BAD_SHAPE,
                 "gemm: non-positive dimension");
  B200SD_REQUIRE((h.c0 % 8) == 0 && (h.c1 % 8) == 0 && (h.ld0 % 8) == 0 && (h.c1 == 0 || (h.ld1 % 8) == 0),
                 B200SD_ERR_BAD_SHAPE, "gemm: channel counts / pixel strides must be multiples of 8");
  B200SD_REQUIRE(h.ld0 >= h.c0 && (h.c1 == 0 || h.ld1 >= h.c1), B200SD_ERR_BAD_SHAPE, "gemm: ld < channels");
  B200SD_REQUIRE((h.c1 == 0) == (h.a1 == nullptr), B200SD_ERR_BAD_ARGUMENT, "gemm: a1 and c1 must agree");
  B200SD_REQUIRE(aligned16(h.a0) && aligned16(h.a1) && aligned16(h.w) && aligned16(h.out) && aligned16(h.bias) &&
                     aligned16(h.residual),
                 B200SD_ERR_BAD_ALIGNMENT, "gemm: pointers must be 16-byte aligned");
  B200SD_REQUIRE(h.act >= 0 && h.act <= 4, B200SD_ERR_BAD_ARGUMENT, "gemm: bad act %d", h.act);
  const bool geglu = (h.act == B200SD_ACT_GEGLU);
  const bool f32 = (h.out_dtype == B200SD_F32);
  const bool res = (h.residual != nullptr);
  B200SD_REQUIRE(!geglu || (h.n % 128) == 0, B200SD_ERR_BAD_SHAPE, "gemm: GEGLU needs n %% 128 == 0");
  B200SD_REQUIRE(!geglu || !f32, B200SD_ERR_BAD_ARGUMENT, "gemm: GEGLU output is 16-bit");
  B200SD_REQUIRE(!geglu || (h.rowbias == nullptr && !res), B200SD_ERR_BAD_ARGUMENT,
                 "gemm: GEGLU epilogue takes no rowbias/residual");
  B200SD_REQUIRE(!res || (!f32 && (h.ldr % 8) == 0), B200SD_ERR_BAD_ARGUMENT,
                 "gemm: residual needs 16-bit output and ldr a multiple of 8");
  const int n_out = geglu ? h.n / 2 : h.n;
  const int esz = f32 ? 4 : 2;
  B200SD_REQUIRE(h.ldc >= n_out, B200SD_ERR_BAD_SHAPE, "gemm: ldc (%d) < output columns (%d)", h.ldc, n_out);
  B200SD_REQUIRE((h.ldc * esz) % 16 == 0, B200SD_ERR_BAD_ALIGNMENT,
                 "gemm: output row pitch (ldc = %d) must be a multiple of 16 bytes", h.ldc);
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
  const int gran = geglu ? 128 : (f32 ? 32 : 64);  // accumulator columns per epilogue group
  // bn_hint: 0 = automatic; otherwise tile width, +1000 forces single-CTA tiles, +2000 forces CTA pairs
  int bn_tile = h.bn_hint % 1000;
  const int force = h.bn_hint / 1000;
  const int nkb_total = f_kh * f_kw * (ceil_div(h.c0, kBlockK) + ceil_div(h.c1, kBlockK));
  const int split = h.split_k > 1 ? h.split_k : 1;
  B200SD_REQUIRE(split <= 16 && split <= nkb_total, B200SD_ERR_BAD_ARGUMENT,
                 "gemm: split_k %d must be <= 16 and <= the %d K blocks of the reduction", split, nkb_total);
  // (K slices of a tile are independent work units: the planner sees split x as many M tiles of 1 / split the depth)
  const TilePlan plan = pick_tiles(d.tiles_m * split, h.n, di.sm_count, gran, nkb_total / split);
  d.pair = split > 1 ? 0 : plan.pair;
  if (bn_tile == 0) bn_tile = plan.bn;
  if (force == 1) d.pair = 0;
  if (force == 2) d.pair = 1;
  if (bn_tile > 256 && force != 2) d.pair = 0;  // wide tiles are single-CTA
  if (split > 1) {
    B200SD_REQUIRE(force != 2, B200SD_ERR_BAD_ARGUMENT, "gemm: split_k runs on single-CTA tiles (no pair mode)");
    if (bn_tile > 256) bn_tile = 256;
  }
  B200SD_REQUIRE(bn_tile >= gran && bn_tile <= (d.pair ? 256 : 512) && (bn_tile % gran) == 0 &&
                     (!d.pair || (bn_tile % 32) == 0) && (bn_tile <= 256 || (bn_tile % 64) == 0),
                 B200SD_ERR_BAD_ARGUMENT, "gemm: bad bn_hint %d (multiple of %d, <= 256, or <= 512 in multiples of 64 "
                 "for single-CTA tiles)", bn_tile, gran);
  d.BN = bn_tile;
  d.wide = bn_tile > 256 ? 1 : 0;
  d.tiles_ncol = ceil_div(h.n, d.BN);
  d.div_munits = make_fastdiv(static_cast<uint32_t>(d.pair ? (d.tiles_m + 1) / 2 : d.tiles_m));
  d.split_k = split;
  d.b_static = h.weights_static != 0 ? 1 : 0;
  d.div_split = make_fastdiv(static_cast<uint32_t>(split));
  d.div_tw = make_fastdiv(static_cast<uint32_t>(d.tiles_w));
  d.div_twth = make_fastdiv(static_cast<uint32_t>(d.tiles_w * d.tiles_h));
  d.N = h.n;
  d.nimg = h.nimg;
  d.H = h.h;
  d.W = h.w_;
  d.kw_taps = f_kw;
  d.taps = f_kh * f_kw;
  d.off_h = f_offh;
  d.off_w = f_offw;
  d.stride = f_stride;
  d.cblk0 = ceil_div(h.c0, kBlockK);
  d.cblk1 = ceil_div(h.c1, kBlockK);
  const uint32_t stage_bytes = kABytes + (d.pair ? d.BN / 2 : d.BN) * kBlockK * 2;
  // residual prefetch (a second staging tile per epilogue warpgroup) for short reductions, where the epilogue sets
  // the pace; long reductions keep the shared memory for mainloop stages (measured: they lose with fewer stages)
  const int budget0 = di.max_smem_optin - 1024 /*align slack*/ - 512 /*static*/;
  const int num_kb_host = d.taps * (d.cblk0 + d.cblk1);
  d.res_bufs = 1;
  if (h.residual != nullptr) {
    const int st2 = (budget0 - 2 * kEpiStageBytes) / static_cast<int>(stage_bytes);
    if (st2 >= 2 && num_kb_host <= 24 && split == 1) d.res_bufs = 2;  // (the finishing CTA of a split tile is not known ahead)
  }
  const int smem_budget = budget0 - d.res_bufs * kEpiStageBytes;
  d.num_stages = min(kMaxStages, smem_budget / static_cast<int>(stage_bytes));
  B200SD_REQUIRE(d.num_stages >= 2, B200SD_ERR_CUDA, "gemm: not enough shared memory (%d B opt-in)", di.max_smem_optin);
  d.bias = h.bias;
  d.rowbias = h.rowbias;
  d.gn_stats = static_cast<unsigned long long*>(h.gn_stats);
  if (h.gn_stats != nullptr) {
    B200SD_REQUIRE(!geglu && !f32, B200SD_ERR_BAD_ARGUMENT, "gemm: fused GroupNorm statistics need a 16-bit, non-GEGLU output");
    B200SD_REQUIRE(h.gn_groups > 0 && (h.n % h.gn_groups) == 0 && ((h.n / h.gn_groups) % 2) == 0, B200SD_ERR_BAD_SHAPE,
                   "gemm: %d output channels / %d groups must be a whole, even number of channels", h.n, h.gn_groups);
    if (h.gn_rows_per_img > 0)
      B200SD_REQUIRE(h.nimg == 1 && h.h == 1 && (h.gn_rows_per_img % 32) == 0, B200SD_ERR_BAD_SHAPE,
                     "gemm: gn_rows_per_img needs the flattened geometry and a multiple of 32 rows per image");
    else
      B200SD_REQUIRE(d.bw * d.bh >= 32, B200SD_ERR_BAD_SHAPE,
                     "gemm: fused GroupNorm statistics need at least 32 pixels of an image per tile (%d x %d)", d.bw, d.bh);
    d.gn_groups = h.gn_groups;
    d.gn_cpg = h.n / h.gn_groups;
    d.gn_rows_per_img = h.gn_rows_per_img;
  }
  d.ld_rowbias = h.ld_rowbias;
  d.act = h.act;
  d.alpha = h.alpha;
  if (h.ln_in_stats != nullptr) {
    B200SD_REQUIRE(h.ln_colsum != nullptr && h.ksize == 1 && !custom && h.c1 == 0 && h.rowbias == nullptr &&
                       h.alpha == 1.f && aligned16(h.ln_in_stats) && aligned16(h.ln_colsum),
                   B200SD_ERR_BAD_ARGUMENT,
                   "gemm: a folded LayerNorm input needs ln_colsum, a plain linear / 1x1 layer over one source, no "
                   "rowbias and alpha = 1");
    d.ln_in = static_cast<const unsigned long long*>(h.ln_in_stats);
    d.ln_colsum = h.ln_colsum;
    d.ln_inv_k = 1.f / static_cast<float>(h.c0);
    d.ln_eps = h.ln_eps;
  }
  if (h.ln_out_stats != nullptr) {
    B200SD_REQUIRE(!f32 && aligned16(h.ln_out_stats), B200SD_ERR_BAD_ARGUMENT,
                   "gemm: row statistics (ln_out_stats) need a 16-bit output");
    d.ln_out = static_cast<unsigned long long*>(h.ln_out_stats);
  }
  {
    // split-K workspace: [tiles] int32 counters (padded to 256 B), then [tiles][split][128][BN] fp32 slices
    const size_t ntiles = static_cast<size_t>(d.tiles_m) * d.tiles_ncol;
    const size_t cnt_bytes = (ntiles * sizeof(int) + 255) & ~static_cast<size_t>(255);
    const size_t need = split > 1 ? cnt_bytes + ntiles * split * kBlockM * d.BN * sizeof(float) : 0;
    if (ws_query != nullptr) {
      *ws_query = need;
      return B200SD_OK;
    }
    if (split > 1) {
      B200SD_REQUIRE(h.workspace != nullptr && static_cast<size_t>(h.workspace_bytes) >= need &&
                         (reinterpret_cast<uintptr_t>(h.workspace) & 255u) == 0,
                     B200SD_ERR_BAD_ARGUMENT,
                     "gemm: split_k = %d needs a zero-initialised, 256-byte aligned workspace of %zu bytes "
                     "(b200sd_gemm_workspace_bytes); got %lld",
                     split, need, static_cast<long long>(h.workspace_bytes));
      d.cnt = static_cast<int*>(h.workspace);
      d.ws = reinterpret_cast<float*>(static_cast<char*>(h.workspace) + cnt_bytes);
    }
  }

  // A: the input feature map [nimg, h * stride, w * stride, c]; a stride-2 box spans 2 bw x 2 bh input pixels and the
  // TMA unit picks every other one (element strides), so the tile in shared memory is the same 128 x 64 as for stride 1
  const uint32_t sA = static_cast<uint32_t>(f_stride);
  const uint32_t boxA[4] = {kBlockK, static_cast<uint32_t>(d.bw) * sA, static_cast<uint32_t>(d.bh) * sA,
                            static_cast<uint32_t>(d.bn)};
  const uint32_t estrA[4] = {1, sA, sA, 1};
  auto in_map = [&](CUtensorMap* tm, const void* ptr, int c, int ld) -> int {
    const uint64_t dims[4] = {static_cast<uint64_t>(c), static_cast<uint64_t>(h.w_) * sA, static_cast<uint64_t>(h.h) * sA,
                              static_cast<uint64_t>(h.nimg)};
    const uint64_t pix = static_cast<uint64_t>(ld) * 2;
    const uint64_t st[3] = {pix, pix * dims[1], pix * dims[1] * dims[2]};
    return encode_tmap_act(tm, ptr, 4, dims, st, boxA, f_stride > 1 ? estrA : nullptr);
  };
  // O / residual: [nimg, h, w, n_out] with pixel stride ld - or explicit pixel / row / image strides (elements) when the
  // output pixels are interleaved into a larger tensor (the four phases of a folded 2x upsample + conv)
  auto out_map = [&](CUtensorMap* tm, const void* ptr, int c, int ld, int elem_bytes, const uint32_t* box,
                     bool as_f32, bool strided) -> int {
    const uint64_t dims[4] = {static_cast<uint64_t>(c), static_cast<uint64_t>(h.w_), static_cast<uint64_t>(h.h),
                              static_cast<uint64_t>(h.nimg)};
    const uint64_t pix = static_cast<uint64_t>(strided ? h.out_pix_stride : ld) * elem_bytes;
    const uint64_t row = strided ? static_cast<uint64_t>(h.out_row_stride) * elem_bytes : pix * h.w_;
    const uint64_t img = strided ? static_cast<uint64_t>(h.out_img_stride) * elem_bytes : row * h.h;
    const uint64_t st[3] = {pix, row, img};
    return as_f32 ? encode_tmap_f32(tm, ptr, 4, dims, st, box) : encode_tmap_act(tm, ptr, 4, dims, st, box);
  };
  const bool out_strided = h.out_pix_stride != 0;
  if (out_strided)
    B200SD_REQUIRE(h.out_pix_stride >= h.ldc && h.out_row_stride >= h.out_pix_stride && h.out_img_stride >= h.out_row_stride &&
                       (h.out_pix_stride * esz) % 16 == 0 && (h.out_row_stride * esz) % 16 == 0 &&
                       (h.out_img_stride * esz) % 16 == 0 && !res,
                   B200SD_ERR_BAD_ARGUMENT, "gemm: bad output strides (16-byte multiples, pix <= row <= img, no residual)");
  if (int e = in_map(&d.tmA0, h.a0, h.c0, h.ld0)) return e;
  if (h.c1 > 0)
    if (int e = in_map(&d.tmA1, h.a1, h.c1, h.ld1)) return e;
  {
    const uint64_t ktot = static_cast<uint64_t>(d.taps) * (d.cblk0 + d.cblk1) * kBlockK;
    const uint64_t dims[2] = {ktot, static_cast<uint64_t>(h.n)};
    const uint64_t st[1] = {ktot * 2};
    const uint32_t box[2] = {kBlockK, static_cast<uint32_t>((d.pair || d.wide) ? d.BN / 2 : d.BN)};
    if (int e = encode_tmap_act(&d.tmB, h.w, 2, dims, st, box)) return e;
  }
  {
    const uint32_t boxO[4] = {static_cast<uint32_t>(f32 ? 32 : 64), static_cast<uint32_t>(d.bw),
                              static_cast<uint32_t>(d.bh), static_cast<uint32_t>(d.bn)};
    if (int e = out_map(&d.tmO, h.out, n_out, h.ldc, esz, boxO, f32, out_strided)) return e;
    if (res)
      if (int e = out_map(&d.tmR, h.residual, n_out, h.ldr, 2, boxO, false, false)) return e;
  }

  const int smem_bytes = d.num_stages * stage_bytes + d.res_bufs * kEpiStageBytes + 1024;
  const int num_tiles = (d.pair ? (d.tiles_m + 1) / 2 : d.tiles_m) * d.tiles_ncol * split;  // work units
  const int grid = d.pair ? 2 * min(num_tiles, di.sm_count / 2) : min(num_tiles, di.sm_count);
#ifdef B200SD_DIAG
  d.diag = getenv("B200SD_GEMM_DIAG") ? atoi(getenv("B200SD_GEMM_DIAG")) : 0;
#endif
  {
    // column-fastest runs whenever there are several column tiles (see GemmDev::col_fast).  Measured inside the SD1.5
    // step at batch 8: 65536 x 1280 -> 320 (168 MB of activations) 83 -> 73 us, 16384 x 2560 -> 640 -6 %,
    // 65536 x 320 -> 960 -7 %, the 64x64-level convs -1..2 %, GEGLU 320 -> 2560 +1 %; all GEMMs 10.36 -> 10.23 ms.
    const char* sw = getenv("B200SD_GEMM_COLFAST");  // "0": M-fastest strided order everywhere (A/B measurements)
    d.col_fast = (d.tiles_ncol > 1 && split == 1 && !(sw != nullptr && sw[0] == '0')) ? 1 : 0;
    const int ctas = d.pair ? grid / 2 : grid;
    d.div_ncol = make_fastdiv(static_cast<uint32_t>(d.tiles_ncol));
    d.div_ctas = make_fastdiv(static_cast<uint32_t>(ctas));
    d.run_q = num_tiles / ctas;
    d.run_rem = num_tiles % ctas;
  }
  GemmLaunch L;
  memset(&L, 0, sizeof(L));
  L.d = d;
  L.grid = grid;
  L.smem_bytes = smem_bytes;
  L.variant = geglu ? 3 : (f32 ? 2 : (res ? 1 : 0));
  {
    const char* sw = getenv("B200SD_GEMM_SIMPLE");  // "0": always the general epilogue (A/B measurements)
    L.simple = (!(sw != nullptr && sw[0] == '0') && !f32 && split == 1 && h.alpha == 1.f && h.rowbias == nullptr &&
                h.ln_in_stats == nullptr && h.ln_out_stats == nullptr && h.gn_stats == nullptr &&
                (h.act == B200SD_ACT_NONE || h.act == B200SD_ACT_GEGLU))
                   ? 1
                   : 0;
  }
  launch_cache_put(0, hp, sizeof(*hp), &L, sizeof(L));
  return launch_cached(L, di.max_smem_optin, stream);
}

extern "C" int b200sd_gemm(const b200sd_gemm_params* hp, b200sd_stream stream_) { return gemm_impl(hp, stream_, nullptr); }

extern "C" size_t b200sd_gemm_workspace_bytes(const b200sd_gemm_params* hp) {
  size_t need = 0;
  if (gemm_impl(hp, nullptr, &need) != B200SD_OK) return 0;
  return need;
}
