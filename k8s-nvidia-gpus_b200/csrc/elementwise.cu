// Small fused, bandwidth-bound kernels around the denoise loop: UNet-input preparation, sinusoidal timestep
// embedding, classifier-free-guidance + scheduler step (DDIM / Euler / PNDM-PLMS), CLIP embedding lookup, image
// denormalisation.  SURVEY §2.1 K10-K12, K17.  (K13 nearest-2x upsample and K2 stride-2 conv live inside the GEMM
// kernel's addressing - gemm.cu - and have no kernels of their own.)
//
// The timestep-dependent kernels read a device-resident step table indexed by a device step counter, so a single
// captured CUDA graph of "one denoise step" is replayed for every step with no host round trip.
#include "common.cuh"
#include "host.cuh"

namespace b200sd {

static constexpr int kEwThreads = 256;

__global__ void prep_unet_input_kernel(const float4* __restrict__ latents, uint4* __restrict__ out,
                                       const float* __restrict__ table, const int* __restrict__ counter, int npix,
                                       int cfg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const float s = table[counter[0] * B200SD_ST_WIDTH + B200SD_ST_IN_SCALE];
  const float4 v = latents[i];
  uint4 u;
  u.x = pack_act2(v.x * s, v.y * s);
  u.y = pack_act2(v.z * s, v.w * s);
  u.z = 0u;
  u.w = 0u;
  out[i] = u;
  if (cfg) out[npix + i] = u;
}

__global__ void timestep_embedding_kernel(act_t* __restrict__ out, const float* __restrict__ table,
                                          const int* __restrict__ counter, int rows, int dim) {
  const int half = dim >> 1;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * half) return;
  const int row = i / half, j = i - row * half;
  const float t = table[counter[0] * B200SD_ST_WIDTH + B200SD_ST_TIMESTEP];
  // diffusers get_timestep_embedding(flip_sin_to_cos=True, downscale_freq_shift=0): [cos | sin]
  const float freq = expf(-9.210340371976184f * static_cast<float>(j) / static_cast<float>(half));
  const float a = t * freq;
  out[static_cast<size_t>(row) * dim + j] = to_act(cosf(a));
  out[static_cast<size_t>(row) * dim + half + j] = to_act(sinf(a));
}

// one thread per latent pixel (4 channels)
__global__ void cfg_sched_kernel(const float* __restrict__ eps, float4* __restrict__ latents,
                                 float4* __restrict__ history, float4* __restrict__ cur_sample,
                                 const float* __restrict__ table, int* __restrict__ counter, int npix, int ld_eps,
                                 int cfg, float gs) {
  const int step = counter[0];
  const float* T = table + step * B200SD_ST_WIDTH;
  if (T[B200SD_ST_GUIDANCE] != 0.f) gs = T[B200SD_ST_GUIDANCE];  // per-call guidance travels in the step table
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < npix) {
    const float* eu = eps + static_cast<size_t>(i) * ld_eps;
    float e[4];
    if (cfg) {
      const float* ec = eps + static_cast<size_t>(npix + i) * ld_eps;
#pragma unroll
      for (int c = 0; c < 4; ++c) e[c] = eu[c] + gs * (ec[c] - eu[c]);
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c) e[c] = eu[c];
    }
    float4 xv = latents[i];
    float x[4] = {xv.x, xv.y, xv.z, xv.w};
    const int kind = static_cast<int>(T[B200SD_ST_KIND]);
    if (kind == 0) {  // DDIM, eta = 0
      const float s1mat = T[B200SD_ST_C0], sat = T[B200SD_ST_C1], sap = T[B200SD_ST_C2], s1map = T[B200SD_ST_C3];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float x0 = (x[c] - s1mat * e[c]) / sat;
        x[c] = sap * x0 + s1map * e[c];
      }
    } else if (kind == 1) {  // Euler discrete, no churn
      const float dt = T[B200SD_ST_C0], sigma = T[B200SD_ST_C1];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float x0 = x[c] - sigma * e[c];
        const float der = (x[c] - x0) / sigma;
        x[c] = x[c] + der * dt;
      }
    } else {  // PNDM / PLMS
      const int mode = static_cast<int>(T[B200SD_ST_PNDM_MODE]);
      float4* h0 = history;
      float4* h1 = history + npix;
      float4* h2 = history + 2 * static_cast<size_t>(npix);
      float4* h3 = history + 3 * static_cast<size_t>(npix);
      if (T[B200SD_ST_PUSH] != 0.f) {
        h3[i] = h2[i];
        h2[i] = h1[i];
        h1[i] = h0[i];
        h0[i] = make_float4(e[0], e[1], e[2], e[3]);
      }
      float m[4];
      if (mode == 1) {
        cur_sample[i] = xv;
#pragma unroll
        for (int c = 0; c < 4; ++c) m[c] = e[c];
      } else if (mode == 2) {
        const float4 hv = h0[i];
        const float hh[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) m[c] = (e[c] + hh[c]) / 2.f;
        const float4 cs = cur_sample[i];
        x[0] = cs.x; x[1] = cs.y; x[2] = cs.z; x[3] = cs.w;
      } else {
        const float w0 = T[B200SD_ST_W0], w1 = T[B200SD_ST_W0 + 1], w2 = T[B200SD_ST_W0 + 2], w3 = T[B200SD_ST_W0 + 3];
        const float4 a = h0[i], b = h1[i], cc = h2[i], d = h3[i];
        const float ha[4] = {a.x, a.y, a.z, a.w}, hb[4] = {b.x, b.y, b.z, b.w}, hc[4] = {cc.x, cc.y, cc.z, cc.w},
                    hd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float acc = w0 * ha[c];
          if (w1 != 0.f) acc += w1 * hb[c];
          if (w2 != 0.f) acc += w2 * hc[c];
          if (w3 != 0.f) acc += w3 * hd[c];
          m[c] = acc;
        }
      }
      const float sc = T[B200SD_ST_C0], mc = T[B200SD_ST_C1];
#pragma unroll
      for (int c = 0; c < 4; ++c) x[c] = sc * x[c] - mc * m[c];
    }
    latents[i] = make_float4(x[0], x[1], x[2], x[3]);
  }
  // last CTA to finish advances the step counter (every CTA has read counter[0] by then)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const int ticket = atomicAdd(&counter[1], 1);
    if (ticket == static_cast<int>(gridDim.x) - 1) {
      counter[1] = 0;
      counter[0] = step + 1;
      __threadfence();
    }
  }
}

__global__ void clip_embed_kernel(const int* __restrict__ ids, const act_t* __restrict__ tok,
                                  const act_t* __restrict__ pos, act_t* __restrict__ out, int rows, int seq, int c,
                                  int vocab) {
  const int octs = c >> 3;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * octs) return;
  const int row = i / octs, oct = i - row * octs;
  int id = ids[row];
  id = min(max(id, 0), vocab - 1);
  const uint4 a = __ldg(reinterpret_cast<const uint4*>(tok + static_cast<size_t>(id) * c) + oct);
  const uint4 b = __ldg(reinterpret_cast<const uint4*>(pos + static_cast<size_t>(row % seq) * c) + oct);
  const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 fa = unpack_act2(av[k]), fb = unpack_act2(bv[k]);
    o[k] = pack_act2(fa.x + fb.x, fa.y + fb.y);
  }
  reinterpret_cast<uint4*>(out + static_cast<size_t>(row) * c)[oct] = make_uint4(o[0], o[1], o[2], o[3]);
}

__global__ void denorm_u8_kernel(const float* __restrict__ x, uint8_t* __restrict__ out, long long npix, int ldx) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= npix) return;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float v = x[i * ldx + c] / 2.f + 0.5f;
    v = fminf(fmaxf(v, 0.f), 1.f);
    out[i * 3 + c] = static_cast<uint8_t>(rintf(v * 255.f));  // round-half-even, like numpy.round
  }
}

__global__ void latent_to_act_kernel(const float4* __restrict__ latents, uint4* __restrict__ out, int npix,
                                      float scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const float4 v = latents[i];
  uint4 u;
  u.x = pack_act2(v.x * scale, v.y * scale);
  u.y = pack_act2(v.z * scale, v.w * scale);
  u.z = 0u;
  u.w = 0u;
  out[i] = u;
}

}  // namespace b200sd

using namespace b200sd;

#define EW_STREAM static_cast<cudaStream_t>(stream_)

extern "C" int b200sd_prep_unet_input(const b200sd_prep_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->latents && p->out && p->step_table && p->step_counter && p->b > 0 && p->hw > 0,
                 B200SD_ERR_BAD_ARGUMENT, "prep_unet_input: bad arguments");
  const int npix = p->b * p->hw;
  prep_unet_input_kernel<<<ceil_div(npix, kEwThreads), kEwThreads, 0, EW_STREAM>>>(
      reinterpret_cast<const float4*>(p->latents), reinterpret_cast<uint4*>(p->out), p->step_table, p->step_counter,
      npix, p->cfg);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_timestep_embedding(const b200sd_temb_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->out && p->step_table && p->step_counter && p->rows > 0 && p->dim > 0 && (p->dim % 2) == 0,
                 B200SD_ERR_BAD_ARGUMENT, "timestep_embedding: bad arguments");
  const int n = p->rows * (p->dim / 2);
  timestep_embedding_kernel<<<ceil_div(n, kEwThreads), kEwThreads, 0, EW_STREAM>>>(
      static_cast<act_t*>(p->out), p->step_table, p->step_counter, p->rows, p->dim);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_cfg_scheduler_step(const b200sd_sched_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->eps && p->latents && p->step_table && p->step_counter && p->b > 0 && p->hw > 0 &&
                     p->ld_eps >= 4,
                 B200SD_ERR_BAD_ARGUMENT, "cfg_scheduler_step: bad arguments");
  const int npix = p->b * p->hw;
  cfg_sched_kernel<<<ceil_div(npix, kEwThreads), kEwThreads, 0, EW_STREAM>>>(
      p->eps, reinterpret_cast<float4*>(p->latents), reinterpret_cast<float4*>(p->history),
      reinterpret_cast<float4*>(p->cur_sample), p->step_table, p->step_counter, npix, p->ld_eps, p->cfg,
      p->guidance_scale);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_clip_embed(const b200sd_embed_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->ids && p->tok && p->pos && p->out && p->rows > 0 && p->seq > 0 && p->c > 0 &&
                     (p->c % 8) == 0 && p->vocab > 0,
                 B200SD_ERR_BAD_ARGUMENT, "clip_embed: bad arguments");
  const int n = p->rows * (p->c / 8);
  clip_embed_kernel<<<ceil_div(n, kEwThreads), kEwThreads, 0, EW_STREAM>>>(
      p->ids, static_cast<const act_t*>(p->tok), static_cast<const act_t*>(p->pos), static_cast<act_t*>(p->out), p->rows,
      p->seq, p->c, p->vocab);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_denorm_u8(const b200sd_denorm_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->x && p->out && p->npix > 0 && p->ldx >= 3, B200SD_ERR_BAD_ARGUMENT,
                 "denorm_u8: bad arguments");
  const long long grid = (p->npix + kEwThreads - 1) / kEwThreads;
  denorm_u8_kernel<<<static_cast<unsigned>(grid), kEwThreads, 0, EW_STREAM>>>(p->x, p->out, p->npix, p->ldx);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_latent_to_act(const b200sd_latent_cast_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->latents && p->out && p->b > 0 && p->hw > 0, B200SD_ERR_BAD_ARGUMENT,
                 "latent_to_act: bad arguments");
  const int npix = p->b * p->hw;
  latent_to_act_kernel<<<ceil_div(npix, kEwThreads), kEwThreads, 0, EW_STREAM>>>(
      reinterpret_cast<const float4*>(p->latents), reinterpret_cast<uint4*>(p->out), npix, p->scale);
  B200SD_CUDA_CHECK(cudaGetLastError());
  return B200SD_OK;
}

extern "C" int b200sd_memset_zero(const b200sd_memset_params* p, b200sd_stream stream_) {
  B200SD_REQUIRE(p && p->ptr && p->bytes >= 0, B200SD_ERR_BAD_ARGUMENT, "memset_zero: bad arguments");
  B200SD_CUDA_CHECK(cudaMemsetAsync(p->ptr, 0, static_cast<size_t>(p->bytes), EW_STREAM));
  return B200SD_OK;
}
