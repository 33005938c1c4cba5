// Host-side helpers shared by the C-ABI entry points: error reporting, device checks, TMA descriptor encoding.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200sd.h"

namespace b200sd {

char* last_error_buf();  // thread-local, defined in capi.cu
int set_error(int code, const char* fmt, ...);

#define B200SD_CUDA_CHECK(expr)                                                                     \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess)                                                                          \
      return b200sd::set_error(B200SD_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                               __LINE__);                                                           \
  } while (0)

#define B200SD_REQUIRE(cond, code, ...)                        \
  do {                                                         \
    if (!(cond)) return b200sd::set_error(code, __VA_ARGS__); \
  } while (0)

struct DeviceInfo {
  int sm_count;
  int cc_major, cc_minor;
  int max_smem_optin;
};
// Cached per current device; returns error if not sm_100.
int get_device_info(DeviceInfo* out);

// cuTensorMapEncodeTiled via the runtime's driver entry point (no -lcuda link dependency).
// dims/strides innermost first; strides[i] = byte stride of dim i+1 (rank-1 entries); act_t elements; 128B swizzle.
// elem_strides (optional, per dim): traversal step inside the box - the TMA unit then loads every elem_strides[i]-th
// element, ceil(box[i] / elem_strides[i]) of them (the stride-2 convolutions read every other pixel this way).
int encode_tmap_act(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box, const uint32_t* elem_strides = nullptr);
// same, fp32 elements (epilogue output maps)
int encode_tmap_f32(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box);

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE property of a kernel: set once per (current device,
// kernel) under a lock, so a second GPU in the same process - or two host threads racing the first call - never
// launch with the default 48 KB limit.
int ensure_dyn_smem(const void* kernel, int bytes);

// Launch-descriptor cache: building a launch costs 3-5 cuTensorMapEncodeTiled calls plus the tile
// plan; a plan replays the same parameter blocks over and over, so the finished device-side descriptor block is kept,
// keyed by (device, kernel family, the caller's parameter bytes).  Mutex-guarded, bounded (cleared when full).
bool launch_cache_get(int family, const void* params, size_t params_bytes, void* out, size_t out_bytes);
void launch_cache_put(int family, const void* params, size_t params_bytes, const void* val, size_t val_bytes);

// attention_cross.cu: persistent short-KV attention (nk <= 128, not causal).  B200SD_OK after launching, > 0 when the
// shape does not fit that kernel (the caller falls through to the general flash kernel), < 0 on error.
int launch_cross_attention(const b200sd_attn_params& h, const DeviceInfo& di, cudaStream_t stream);

// Programmatic dependent launch per kernel family; B200SD_PDL is a bit mask (1 GEMM, 2 attention, 4 norms; 0 = off;
// default 1), read once.  Every kernel family issues griddepcontrol.launch_dependents early either way, so a GEMM that
// follows a norm or attention kernel starts its set-up on the SMs that kernel has already left.
enum { kPdlGemm = 1, kPdlAttention = 2, kPdlNorm = 4 };
bool pdl_enabled(int family);

// kernel<<<grid, block, smem, stream>>>(args...) with programmatic dependent launch allowed: every kernel launched
// through this calls pdl_wait() (common.cuh) before its first access to global memory.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(int family, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                     Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled(family) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace b200sd
