// Bandwidth-bound normalisation kernels: GroupNorm(+SiLU) over a virtual two-source concat, LayerNorm, row softmax.
// All statistics are fp32; inputs/outputs act_t, 16-byte vectorised, coalesced along the NHWC channel axis.
// SURVEY §2.1 K7, K8, K9, K14.
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kNormThreads = 256;

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const float2 a = unpack_act2(u.x), b = unpack_act2(u.y), c = unpack_act2(u.z), d = unpack_act2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ void load8(const act_t* p, float (&f)[8]) {
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
  const float2 a = unpack_act2(u.x), b = unpack_act2(u.y), c = unpack_act2(u.z), d = unpack_act2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ void store8(act_t* p, const float (&f)[8]) {
  uint4 u;
  u.x = pack_act2(f[0], f[1]);
  u.y = pack_act2(f[2], f[3]);
  u.z = pack_act2(f[4], f[5]);
  u.w = pack_act2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// ---------------------------------------------------------------------------------------------- GroupNorm: stats
// grid (chunks, nimg).  Each thread owns one 8-channel vector column ("octet") and a pixel stream, keeping 8 sums and
// 8 sums of squares in registers.  The CTA's partial sums are combined in shared memory by a FIXED-order fp32 tree
// (octet sums of every pixel stream -> one thread per (group, statistic)), converted once per (group, statistic, CTA)
// to 64-bit fixed point (common.cuh, fixed_split) and added to the global slot as integers: the totals are
// bit-identical from run to run whatever the CTA arrival order, and - pixels per CTA being a function of the
// per-image shape only - whatever batch the image rides in.  (An earlier version converted per thread and used 64-bit
// shared-memory atomics: 8.7 instructions per element, most of them in that tail.)
__global__ void __launch_bounds__(kNormThreads) gn_stats_kernel(const act_t* __restrict__ x0,
                                                                const act_t* __restrict__ x1,
                                                                unsigned long long* __restrict__ stats, int hw, int c0,
                                                                int c1, int groups, int pix_per_cta) {
  extern __shared__ float gred[];  // [streams][octs][16]: {8 sums, 8 sums of squares} of a (pixel stream, octet)
  const int n = blockIdx.y;
  const int C = c0 + c1;
  const int octs = C >> 3;
  const int cpg = C / groups;
  const int p_begin = blockIdx.x * pix_per_cta;
  const int p_end = min(hw, p_begin + pix_per_cta);
  const int Wd = min(octs, kNormThreads);
  const int S = kNormThreads / Wd;
  const int stream = threadIdx.x / Wd;
  pdl_launch();
  pdl_wait();  // programmatic dependent launch (common.cuh): nothing of the step is touched before this
  if (stream < S) {
    for (int oct = threadIdx.x % Wd; oct < octs; oct += Wd) {
      const int ch0 = oct * 8;
      const act_t* src;
      int ld;
      if (ch0 < c0) {
        src = x0 + static_cast<size_t>(n) * hw * c0 + ch0;
        ld = c0;
      } else {
        src = x1 + static_cast<size_t>(n) * hw * c1 + (ch0 - c0);
        ld = c1;
      }
      float s[8], q[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) s[e] = q[e] = 0.f;
      int p = p_begin + stream;
      for (; p + 7 * S < p_end; p += 8 * S) {
        uint4 raw[8];  // eight pixels in flight, kept packed until they are used
#pragma unroll
        for (int u = 0; u < 8; ++u) raw[u] = __ldg(reinterpret_cast<const uint4*>(src + static_cast<size_t>(p + u * S) * ld));
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          float f[8];
          unpack8(raw[u], f);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            s[e] += f[e];
            q[e] = fmaf(f[e], f[e], q[e]);
          }
        }
      }
      for (; p < p_end; p += S) {
        float f[8];
        load8(src + static_cast<size_t>(p) * ld, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          s[e] += f[e];
          q[e] = fmaf(f[e], f[e], q[e]);
        }
      }
      float4* dst = reinterpret_cast<float4*>(gred + (static_cast<size_t>(stream) * octs + oct) * 16);
      dst[0] = make_float4(s[0], s[1], s[2], s[3]);
      dst[1] = make_float4(s[4], s[5], s[6], s[7]);
      dst[2] = make_float4(q[0], q[1], q[2], q[3]);
      dst[3] = make_float4(q[4], q[5], q[6], q[7]);
    }
  }
  __syncthreads();
  // one thread per (group, statistic): the group's channels x the pixel streams, always in the same order
  for (int i = threadIdx.x; i < 2 * groups; i += blockDim.x) {
    const int g = i >> 1, which = i & 1;
    float acc = 0.f;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) {
      const float* col = gred + (c >> 3) * 16 + which * 8 + (c & 7);
      for (int st = 0; st < S; ++st) acc += col[static_cast<size_t>(st) * octs * 16];
    }
    long long H, L;
    fixed_split(acc, H, L);
    unsigned long long* slot = stats + (static_cast<size_t>(n) * groups + g) * kGnSlotWords64 + which * 2;
    atomicAdd(slot, static_cast<unsigned long long>(H));
    atomicAdd(slot + 1, static_cast<unsigned long long>(L));
  }
}

// ---------------------------------------------------------------------------------------------- GroupNorm: apply
// Same thread -> (octet, pixel stream) mapping as the statistics pass: a thread normalises ONE 8-channel column, so its
// 8 scales and 8 shifts live in registers for the whole walk (no per-element shared-memory lookups) and the inner
// loop is load16 -> 8 fma (+SiLU) -> store16 with four pixels in flight.  The (mean, rstd) of the image's groups are
// finalised once per CTA (one thread per group, in double from the integer totals) into shared memory.  kSilu is a
// template parameter: as a run-time value the compiler evaluated both SiLU forms per element and selected.
template <int kSilu>
__global__ void __launch_bounds__(kNormThreads, 4) gn_apply_kernel(const act_t* __restrict__ x0,
                                                                const act_t* __restrict__ x1,
                                                                const unsigned long long* __restrict__ stats,
                                                                const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, act_t* __restrict__ out,
                                                                int hw, int c0, int c1, int groups, int pix_per_cta,
                                                                float eps) {
  extern __shared__ float2 gmr[];  // [groups] (mean, rstd)
  pdl_launch();
  pdl_wait();
  const int n = blockIdx.y;
  const int C = c0 + c1;
  const int octs = C >> 3;
  const int cpg = C / groups;
  const double inv_cnt = 1.0 / (static_cast<double>(hw) * static_cast<double>(cpg));
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    // exact integer totals -> mean / variance in double (no E[x^2] - mean^2 cancellation in fp32)
    const unsigned long long* sl = stats + (static_cast<size_t>(n) * groups + g) * kGnSlotWords64;
    const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(sl);
    const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(sl + 2);
    const unsigned long long sw[2] = {a.x, a.y}, qw[2] = {b.x, b.y};
    const double m = fixed_value(sw) * inv_cnt;
    const double var = fmax(fixed_value(qw) * inv_cnt - m * m, 0.0);
    gmr[g] = make_float2(static_cast<float>(m), rsqrtf(static_cast<float>(var) + eps));
  }
  __syncthreads();
  const int p_begin = blockIdx.x * pix_per_cta;
  const int p_end = min(hw, p_begin + pix_per_cta);
  const int Wd = min(octs, kNormThreads);
  const int S = kNormThreads / Wd;
  const int stream = threadIdx.x / Wd;
  if (stream >= S) return;
  auto act = [](float y) { return kSilu == 2 ? silu_tanh_f(y) : (kSilu == 1 ? silu_f(y) : y); };
  for (int oct = threadIdx.x % Wd; oct < octs; oct += Wd) {
    const int ch0 = oct * 8;
    float sc[8], sh[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float2 mr = gmr[(ch0 + e) / cpg];
      sc[e] = __ldg(gamma + ch0 + e) * mr.y;
      sh[e] = __ldg(beta + ch0 + e) - mr.x * sc[e];
    }
    const act_t* src;
    int ld;
    if (ch0 < c0) {
      src = x0 + static_cast<size_t>(n) * hw * c0 + ch0;
      ld = c0;
    } else {
      src = x1 + static_cast<size_t>(n) * hw * c1 + (ch0 - c0);
      ld = c1;
    }
    act_t* dst = out + static_cast<size_t>(n) * hw * C + ch0;
    int p = p_begin + stream;
    for (; p + 3 * S < p_end; p += 4 * S) {
      uint4 raw[4];  // four pixels in flight, kept packed until they are used
#pragma unroll
      for (int u = 0; u < 4; ++u) raw[u] = __ldg(reinterpret_cast<const uint4*>(src + static_cast<size_t>(p + u * S) * ld));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8(raw[u], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = act(fmaf(f[e], sc[e], sh[e]));
        store8(dst + static_cast<size_t>(p + u * S) * C, f);
      }
    }
    for (; p < p_end; p += S) {
      float f[8];
      load8(src + static_cast<size_t>(p) * ld, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = act(fmaf(f[e], sc[e], sh[e]));
      store8(dst + static_cast<size_t>(p) * C, f);
    }
  }
}

// the three activation modes of gn_apply_kernel (0 none, 1 SiLU with ex2 + rcp, 2 SiLU in tanh form)
template <typename... Args>
static cudaError_t launch_gn_apply(bool pdl, int mode, dim3 grid, int groups, cudaStream_t s, Args... args) {
  const size_t smem = static_cast<size_t>(groups) * sizeof(float2);
  auto go = [&](auto kernel) -> cudaError_t {
    if (pdl) return launch_pdl(kPdlNorm, kernel, grid, dim3(kNormThreads), smem, s, args...);
    kernel<<<grid, kNormThreads, smem, s>>>(args...);
    return cudaGetLastError();
  };
  if (mode == 2) return go(gn_apply_kernel<2>);
  if (mode == 1) return go(gn_apply_kernel<1>);
  return go(gn_apply_kernel<0>);
}

// ---------------------------------------------------------------------------------------------- LayerNorm
// kLPR lanes share a row (32 / kLPR rows per warp), each lane holding up to kVecs 8-channel vectors in registers:
// several rows' loads are in flight per warp and the two reductions take log2(kLPR) shuffles.  Two-pass mean /
// variance like the fp32 reference.
template <int kLPR, int kVecs>
__global__ void __launch_bounds__(kNormThreads, (kVecs <= 3 ? 5 : kVecs <= 5 ? 4 : 1)) layernorm_kernel(const act_t* __restrict__ x,
                                                                 const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta, act_t* __restrict__ out,
                                                                 int rows, int c, float eps) {
  pdl_launch();
  pdl_wait();
  constexpr int kRowsPerWarp = 32 / kLPR;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane / kLPR, l = lane % kLPR;
  const int row = (blockIdx.x * (kNormThreads / 32) + warp) * kRowsPerWarp + sub;
  const bool row_ok = row < rows;
  const int octs = c >> 3;
  const act_t* xr = x + static_cast<size_t>(row_ok ? row : 0) * c;
  float f[kVecs][8];
  float sum = 0.f;
#pragma unroll
  for (int v = 0; v < kVecs; ++v) {
    const int oct = l + kLPR * v;
    if (oct < octs) {
      load8(xr + oct * 8, f[v]);
#pragma unroll
      for (int e = 0; e < 8; ++e) sum += f[v][e];
    }
  }
#pragma unroll
  for (int o = kLPR / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum / static_cast<float>(c);
  float sq = 0.f;
#pragma unroll
  for (int v = 0; v < kVecs; ++v) {
    const int oct = l + kLPR * v;
    if (oct < octs) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = f[v][e] - mean;
        sq = fmaf(d, d, sq);
      }
    }
  }
#pragma unroll
  for (int o = kLPR / 2; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq / static_cast<float>(c) + eps);
  if (!row_ok) return;
  act_t* orow = out + static_cast<size_t>(row) * c;
#pragma unroll
  for (int v = 0; v < kVecs; ++v) {
    const int oct = l + kLPR * v;
    if (oct < octs) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + oct * 8));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + oct * 8) + 1);
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + oct * 8));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta + oct * 8) + 1);
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float y[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) y[e] = (f[v][e] - mean) * rstd * gg[e] + bb[e];
      store8(orow + oct * 8, y);
    }
  }
}

// ---------------------------------------------------------------------------------------------- row softmax (in place)
// one CTA per row, row held in registers (cols <= 256 * 8 * kVecs).
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

template <int kVecs>
__global__ void __launch_bounds__(kNormThreads) softmax_rows_kernel(const float* __restrict__ x,
                                                                    act_t* __restrict__ out, int cols) {
  __shared__ float red[kNormThreads / 32];
  const float* xr = x + static_cast<size_t>(blockIdx.x) * cols;
  act_t* orow = out + static_cast<size_t>(blockIdx.x) * cols;
  const int octs = cols >> 3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float f[kVecs][8];
  float mx = -INFINITY;
#pragma unroll
  for (int v = 0; v < kVecs; ++v) {
    const int oct = threadIdx.x + kNormThreads * v;
    if (oct < octs) {
      load8f(xr + oct * 8, f[v]);
#pragma unroll
      for (int e = 0; e < 8; ++e) mx = fmaxf(mx, f[v][e]);
    }
  }
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = red[0];
#pragma unroll
  for (int i = 1; i < kNormThreads / 32; ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float sum = 0.f;
#pragma unroll
  for (int v = 0; v < kVecs; ++v) {
    const int oct = threadIdx.x + kNormThreads * v;
    if (oct < octs) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        f[v][e] = __expf(f[v][e] - mx);
        sum += f[v][e];
      }
    }
  }
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int i = 0; i < kNormThreads / 32; ++i) sum += red[i];
  const float inv = 1.f / sum;
#pragma unroll
  for (int v = 0; v < kVecs; ++v) {
    const int oct = threadIdx.x + kNormThreads * v;
    if (oct < octs) {
#pragma unroll
      for (int e = 0; e < 8; ++e) f[v][e] *= inv;
      store8(orow + oct * 8, f[v]);
    }
  }
}

// shared memory of gn_stats_kernel: 16 floats per (pixel stream, octet)
static size_t gn_stats_smem(int C) {
  const int octs = C / 8;
  const int Wd = octs < kNormThreads ? octs : kNormThreads;
  return static_cast<size_t>(kNormThreads / Wd) * octs * 16 * sizeof(float);
}

// SiLU after GroupNorm takes ONE MUFU op (tanh form, common.cuh silu_tanh_f) instead of two (ex2 + rcp): with two, the
// MUFU pipe is 40 % of gn_apply's time.  Measured: gn_apply 0.92 -> 0.83 ms per UNet step, 2.63 -> 2.37 ms per VAE decode;
// 25-step SD1.5 latent error 8.9e-4 vs 8.7e-4, PSNR 61.0 vs 61.1 dB.  B200SD_SILU_TANH=0 selects the two-op form.
static int gn_silu_mode(int silu) {
  if (!silu) return 0;
  const char* sw = getenv("B200SD_SILU_TANH");
  return (sw != nullptr && sw[0] == '0') ? 1 : 2;
}

static int gn_check(const b200sd_groupnorm_params& h, int* pix_per_cta, int* chunks) {
  B200SD_REQUIRE(h.x0 && h.stats && h.nimg > 0 && h.hw > 0 && h.c0 > 0 && h.c1 >= 0 && h.groups > 0,
                 B200SD_ERR_BAD_ARGUMENT, "groupnorm: bad arguments");
  const int C = h.c0 + h.c1;
  B200SD_REQUIRE((h.c0 % 8) == 0 && (h.c1 % 8) == 0, B200SD_ERR_BAD_SHAPE, "groupnorm: channels must be multiples of 8");
  B200SD_REQUIRE((C % h.groups) == 0, B200SD_ERR_BAD_SHAPE, "groupnorm: %d channels not divisible by %d groups", C,
                 h.groups);
  B200SD_REQUIRE(h.groups <= 1024 && C <= 6144, B200SD_ERR_TOO_LARGE, "groupnorm: too many groups/channels");
  B200SD_REQUIRE(h.nimg <= 65535, B200SD_ERR_TOO_LARGE, "groupnorm: at most 65535 images per call (got %d)", h.nimg);
  B200SD_REQUIRE((h.c1 == 0) == (h.x1 == nullptr), B200SD_ERR_BAD_ARGUMENT, "groupnorm: x1 and c1 must agree");
  B200SD_REQUIRE(aligned16(h.x0) && aligned16(h.x1) && aligned16(h.out), B200SD_ERR_BAD_ALIGNMENT,
                 "groupnorm: pointers must be 16-byte aligned");
  // Pixels per CTA: ~64 KB of input, a function of the per-image shape ONLY - never of the batch size - so that an
  // image's partial sums (and with them its pixels) are bit-identical whatever batch it rides in (batch / sharding
  // invariance).
  int ppc = (64 << 10) / (2 * C);  // (measured: 32 KB pieces 0.343 ms, 64 KB 0.303 ms, 128 KB 0.332 ms of statistics kernels per SD1.5 step at batch 8)
  if (ppc < 8) ppc = 8;
  if (ppc > h.hw) ppc = h.hw;
  *pix_per_cta = ppc;
  *chunks = (h.hw + ppc - 1) / ppc;
  return B200SD_OK;
}

// The apply pass is elementwise, so its split is free of the invariance rule above: pixels per CTA are chosen so that
// the whole grid is resident at once (4 CTAs per SM) with equal shares - no second, mostly empty wave (the
// statistics split gave 656 CTAs for 592 places on the 64x64 level at batch 8), and a full machine at small batch.
static int gn_apply_split(const b200sd_groupnorm_params& h, int* pix_per_cta, int* chunks) {
  DeviceInfo di;
  if (int e = get_device_info(&di)) return e;
  // (large tensors keep the statistics pass's ~64 KB pieces, passed in: with many waves of short CTAs the machine
  // balances itself, and one wave of long-lived CTAs measured 8 % slower on the VAE's 512x512 levels)
  if (static_cast<long long>(*chunks) * h.nimg >= 4LL * di.sm_count * 4) return B200SD_OK;
  const int octs = (h.c0 + h.c1) / 8;
  const int streams = kNormThreads / (octs < kNormThreads ? octs : kNormThreads);
  const int per_img = di.sm_count * 4 / h.nimg > 0 ? di.sm_count * 4 / h.nimg : 1;
  int ppc = (h.hw + per_img - 1) / per_img;
  if (ppc < 4 * streams) ppc = 4 * streams;  // at least one full round of loads per thread
  if (ppc > h.hw) ppc = h.hw;
  *pix_per_cta = ppc;
  *chunks = (h.hw + ppc - 1) / ppc;
  return B200SD_OK;
}

}  // namespace b200sd

using namespace b200sd;

extern "C" int b200sd_groupnorm_stats(const b200sd_groupnorm_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "groupnorm_stats: params is NULL");
  int ppc, chunks;
  if (int e = gn_check(*hp, &ppc, &chunks)) return e;
  const b200sd_groupnorm_params& h = *hp;
  dim3 grid(chunks, h.nimg);
  B200SD_CUDA_CHECK(launch_pdl(kPdlNorm, gn_stats_kernel, grid, dim3(kNormThreads), gn_stats_smem(h.c0 + h.c1),
                               static_cast<cudaStream_t>(stream_), static_cast<const act_t*>(h.x0),
                               static_cast<const act_t*>(h.x1), static_cast<unsigned long long*>(h.stats), h.hw, h.c0,
                               h.c1, h.groups, ppc));
  return B200SD_OK;
}

extern "C" int b200sd_groupnorm_apply(const b200sd_groupnorm_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "groupnorm_apply: params is NULL");
  int ppc, chunks;
  if (int e = gn_check(*hp, &ppc, &chunks)) return e;
  const b200sd_groupnorm_params& h = *hp;
  B200SD_REQUIRE(h.out && h.gamma && h.beta, B200SD_ERR_BAD_ARGUMENT, "groupnorm_apply: out/gamma/beta required");
  if (int e = gn_apply_split(h, &ppc, &chunks)) return e;
  dim3 grid(chunks, h.nimg);
  B200SD_CUDA_CHECK(launch_gn_apply(true, gn_silu_mode(h.silu), grid, h.groups, static_cast<cudaStream_t>(stream_),
                                    static_cast<const act_t*>(h.x0), static_cast<const act_t*>(h.x1),
                                    static_cast<const unsigned long long*>(h.stats), h.gamma, h.beta,
                                    static_cast<act_t*>(h.out), h.hw, h.c0, h.c1, h.groups, ppc, h.eps));
  return B200SD_OK;
}

extern "C" size_t b200sd_groupnorm_workspace_bytes(const b200sd_groupnorm_params* p) {
  if (!p || p->nimg <= 0 || p->groups <= 0) return 0;
  return static_cast<size_t>(p->nimg) * p->groups * B200SD_GN_STATS_WORDS * 4;
}

// ---------------------------------------------------------------------------------------------- GroupNorm, no allocation
// Built-in statistics scratch for b200sd_groupnorm_noalloc: lives in the library image, so every GPU that loads
// libb200sd.so carries these 64 KiB until the process exits.  The name below is only a device symbol — host code
// must not treat it as a pointer; the device address is looked up once per device with cudaGetSymbolAddress.
static __device__ __align__(256) float g_gn_noalloc_scratch[B200SD_GN_NOALLOC_MAX_STATS];

namespace b200sd {
static std::mutex g_gn_noalloc_mu;
static float* g_gn_noalloc_ptr[64] = {nullptr};

// Device address of the built-in scratch on the CURRENT device (resolved on first use, then cached per device).
static int gn_noalloc_scratch(float** out) {
  int dev = 0;
  B200SD_CUDA_CHECK(cudaGetDevice(&dev));
  B200SD_REQUIRE(dev >= 0 && dev < 64, B200SD_ERR_CUDA, "groupnorm_noalloc: device index %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_gn_noalloc_mu);
  if (g_gn_noalloc_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    B200SD_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_gn_noalloc_scratch));
    B200SD_REQUIRE(ptr != nullptr, B200SD_ERR_CUDA, "groupnorm_noalloc: scratch symbol has no device address");
    g_gn_noalloc_ptr[dev] = ptr;
  }
  *out = g_gn_noalloc_ptr[dev];
  return B200SD_OK;
}
}  // namespace b200sd

extern "C" int b200sd_groupnorm_noalloc(const b200sd_groupnorm_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "groupnorm_noalloc: params is NULL");
  B200SD_REQUIRE(hp->nimg > 0 && hp->groups > 0, B200SD_ERR_BAD_ARGUMENT, "groupnorm_noalloc: bad arguments");
  const long long need = static_cast<long long>(B200SD_GN_STATS_WORDS) * hp->nimg * hp->groups;  // 4-byte words
  B200SD_REQUIRE(need <= B200SD_GN_NOALLOC_MAX_STATS, B200SD_ERR_TOO_LARGE,
                 "groupnorm_noalloc: nimg * groups * 8 = %lld statistics words exceed the built-in scratch of %d words "
                 "(use b200sd_groupnorm_stats/_apply with caller-provided scratch for larger problems)",
                 need, B200SD_GN_NOALLOC_MAX_STATS);
  b200sd_groupnorm_params h = *hp;  // same kernels as the two-call path; only the scratch pointer differs
  float* scratch = nullptr;
  if (int e = gn_noalloc_scratch(&scratch)) return e;
  h.stats = scratch;
  int ppc, chunks;
  if (int e = gn_check(h, &ppc, &chunks)) return e;
  B200SD_REQUIRE(h.out && h.gamma && h.beta, B200SD_ERR_BAD_ARGUMENT, "groupnorm_noalloc: out/gamma/beta required");
  const int C = h.c0 + h.c1;
  (void)C;
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  // the scratch still holds the previous call's sums: the accumulators of this call must start from zero
  B200SD_CUDA_CHECK(cudaMemsetAsync(h.stats, 0, static_cast<size_t>(need) * sizeof(float), s));
  dim3 grid(chunks, h.nimg);
  gn_stats_kernel<<<grid, kNormThreads, gn_stats_smem(h.c0 + h.c1), s>>>(
      static_cast<const act_t*>(h.x0), static_cast<const act_t*>(h.x1), static_cast<unsigned long long*>(h.stats), h.hw,
      h.c0, h.c1, h.groups, ppc);
  B200SD_CUDA_CHECK(cudaGetLastError());
  if (int e = gn_apply_split(h, &ppc, &chunks)) return e;
  grid = dim3(chunks, h.nimg);
  B200SD_CUDA_CHECK(launch_gn_apply(false, gn_silu_mode(h.silu), grid, h.groups, s, static_cast<const act_t*>(h.x0),
                                    static_cast<const act_t*>(h.x1), static_cast<const unsigned long long*>(h.stats),
                                    h.gamma, h.beta, static_cast<act_t*>(h.out), h.hw, h.c0, h.c1, h.groups, ppc,
                                    h.eps));
  return B200SD_OK;
}

extern "C" int b200sd_layernorm(const b200sd_layernorm_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "layernorm: params is NULL");
  const b200sd_layernorm_params& h = *hp;
  B200SD_REQUIRE(h.x && h.out && h.gamma && h.beta && h.rows > 0 && h.c > 0, B200SD_ERR_BAD_ARGUMENT,
                 "layernorm: bad arguments");
  B200SD_REQUIRE((h.c % 8) == 0 && h.c <= 4096, B200SD_ERR_BAD_SHAPE, "layernorm: c must be a multiple of 8, <= 4096");
  B200SD_REQUIRE(aligned16(h.x) && aligned16(h.out), B200SD_ERR_BAD_ALIGNMENT, "layernorm: 16-byte alignment");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const act_t* x = static_cast<const act_t*>(h.x);
  act_t* o = static_cast<act_t*>(h.out);
  const int octs = h.c / 8;
  const int warps = kNormThreads / 32;
  B200SD_REQUIRE(aligned16(h.gamma) && aligned16(h.beta), B200SD_ERR_BAD_ALIGNMENT, "layernorm: gamma/beta alignment");
  static const bool narrow = [] {  // A/B switch: fewer row vectors per lane -> 48 registers, 5 CTAs per SM without spills
                                   // (0.616 -> 0.561 ms per step; 6 CTAs at 40 registers spilled and gained nothing)
    const char* e = getenv("B200SD_LN_NARROW");
    return !(e != nullptr && e[0] == '0');
  }();
  auto go = [&](auto kernel, int rows_per_cta) -> cudaError_t {
    return launch_pdl(kPdlNorm, kernel, dim3(ceil_div(h.rows, rows_per_cta)), dim3(kNormThreads), 0, s, x, h.gamma,
                      h.beta, o, h.rows, h.c, h.eps);
  };
  if (octs <= 16 * 3 && narrow)  // <= 384 channels: 16 lanes per row, 2 rows per warp
    B200SD_CUDA_CHECK(go(layernorm_kernel<16, 3>, warps * 2));
  else if (octs <= 8 * 5)  // <= 320 channels: 8 lanes per row, 4 rows per warp
    B200SD_CUDA_CHECK(go(layernorm_kernel<8, 5>, warps * 4));
  else if (octs <= 32 * 3 && narrow)  // <= 768
    B200SD_CUDA_CHECK(go(layernorm_kernel<32, 3>, warps));
  else if (octs <= 16 * 5)  // <= 640
    B200SD_CUDA_CHECK(go(layernorm_kernel<16, 5>, warps * 2));
  else if (octs <= 32 * 5)  // <= 1280
    B200SD_CUDA_CHECK(go(layernorm_kernel<32, 5>, warps));
  else
    B200SD_CUDA_CHECK(go(layernorm_kernel<32, 16>, warps));
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_softmax_rows(const b200sd_softmax_params* hp, b200sd_stream stream_) {
  B200SD_REQUIRE(hp != nullptr, B200SD_ERR_BAD_ARGUMENT, "softmax_rows: params is NULL");
  const b200sd_softmax_params& h = *hp;
  B200SD_REQUIRE(h.x && h.out && h.rows > 0 && h.cols > 0, B200SD_ERR_BAD_ARGUMENT, "softmax_rows: bad arguments");
  B200SD_REQUIRE((h.cols % 8) == 0 && h.cols <= kNormThreads * 8 * 8, B200SD_ERR_BAD_SHAPE,
                 "softmax_rows: cols must be a multiple of 8, <= %d", kNormThreads * 8 * 8);
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  act_t* o = static_cast<act_t*>(h.out);
  if (h.cols <= kNormThreads * 8 * 2)
    softmax_rows_kernel<2><<<h.rows, kNormThreads, 0, s>>>(h.x, o, h.cols);
  else
    softmax_rows_kernel<8><<<h.rows, kNormThreads, 0, s>>>(h.x, o, h.cols);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}
