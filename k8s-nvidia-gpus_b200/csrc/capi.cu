// C-ABI plumbing: version, thread-local error string, device query, TMA descriptor encoding.
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include <stdlib.h>
#include <string.h>

#include "host.cuh"

namespace b200sd {

static thread_local char g_err[512] = "";

char* last_error_buf() { return g_err; }

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

static std::mutex g_mu;
static DeviceInfo g_info[64];
static bool g_info_ok[64] = {false};

int get_device_info(DeviceInfo* out) {
  int dev = 0;
  B200SD_CUDA_CHECK(cudaGetDevice(&dev));
  B200SD_REQUIRE(dev >= 0 && dev < 64, B200SD_ERR_CUDA, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_info_ok[dev]) {
    DeviceInfo di;
    B200SD_CUDA_CHECK(cudaDeviceGetAttribute(&di.sm_count, cudaDevAttrMultiProcessorCount, dev));
    B200SD_CUDA_CHECK(cudaDeviceGetAttribute(&di.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    B200SD_CUDA_CHECK(cudaDeviceGetAttribute(&di.cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    B200SD_CUDA_CHECK(cudaDeviceGetAttribute(&di.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    B200SD_REQUIRE(di.cc_major == 10, B200SD_ERR_UNSUPPORTED_ARCH,
                   "b200sd requires an sm_100-class GPU (found sm_%d%d); there is no fallback path", di.cc_major,
                   di.cc_minor);
    g_info[dev] = di;
    g_info_ok[dev] = true;
  }
  *out = g_info[dev];
  return B200SD_OK;
}

static std::set<std::pair<int, const void*>> g_smem_set;

int ensure_dyn_smem(const void* kernel, int bytes) {
  int dev = 0;
  B200SD_CUDA_CHECK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_mu);
  const auto key = std::make_pair(dev, kernel);
  if (g_smem_set.count(key)) return B200SD_OK;
  B200SD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  g_smem_set.insert(key);
  return B200SD_OK;
}

bool pdl_enabled(int family) {
  static const int mask = [] {
    const char* e = getenv("B200SD_PDL");
    return e != nullptr ? atoi(e) : kPdlGemm;  // measured (profiles/README.md): the attribute pays on the GEMMs only
  }();
  return (mask & family) != 0;
}

static std::mutex g_cache_mu;
static std::unordered_map<std::string, std::vector<unsigned char>> g_launch_cache;
static const size_t kLaunchCacheMax = 8192;

static std::string cache_key(int family, const void* params, size_t n) {
  int dev = 0;
  cudaGetDevice(&dev);
  std::string k(reinterpret_cast<const char*>(params), n);
  k.push_back(static_cast<char>(family));
  k.push_back(static_cast<char>(dev));
  return k;
}

bool launch_cache_get(int family, const void* params, size_t params_bytes, void* out, size_t out_bytes) {
  const std::string k = cache_key(family, params, params_bytes);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  auto it = g_launch_cache.find(k);
  if (it == g_launch_cache.end() || it->second.size() != out_bytes) return false;
  memcpy(out, it->second.data(), out_bytes);
  return true;
}

void launch_cache_put(int family, const void* params, size_t params_bytes, const void* val, size_t val_bytes) {
  const std::string k = cache_key(family, params, params_bytes);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  if (g_launch_cache.size() >= kLaunchCacheMax) g_launch_cache.clear();
  const unsigned char* v = static_cast<const unsigned char*>(val);
  g_launch_cache[k] = std::vector<unsigned char>(v, v + val_bytes);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
#ifdef B200SD_ACT_BF16
static const CUtensorMapDataType kTmaDtype = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
static const int kActCode = 0;
#else
static const CUtensorMapDataType kTmaDtype = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
static const int kActCode = 1;
#endif

static int encode_tmap_any(CUtensorMap* tm, CUtensorMapDataType dtype, const void* base, int rank, const uint64_t* dims,
                           const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides) {
  if (g_encode == nullptr) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_encode == nullptr) {
      void* fn = nullptr;
      cudaDriverEntryPointQueryResult qres;
      B200SD_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
      B200SD_REQUIRE(fn != nullptr && qres == cudaDriverEntryPointSuccess, B200SD_ERR_CUDA,
                     "cuTensorMapEncodeTiled not available from the driver");
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    }
  }
  cuuint64_t gdims[5];
  cuuint64_t gstr[5];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = elem_strides ? elem_strides[i] : 1;
    if (i + 1 < rank) gstr[i] = strides_bytes[i];
  }
  for (int i = 0; i + 1 < rank; ++i)
    B200SD_REQUIRE((gstr[i] % 16) == 0, B200SD_ERR_BAD_ALIGNMENT, "TMA stride %d (%llu B) not a multiple of 16", i,
                   (unsigned long long)gstr[i]);
  CUresult r = g_encode(tm, dtype, static_cast<cuuint32_t>(rank), const_cast<void*>(base),
                        gdims, gstr, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200SD_REQUIRE(r == CUDA_SUCCESS, B200SD_ERR_CUDA,
                 "cuTensorMapEncodeTiled failed (CUresult %d; rank %d dims %llu,%llu,%llu,%llu box %u,%u,%u,%u)", (int)r,
                 rank, (unsigned long long)gdims[0], (unsigned long long)(rank > 1 ? gdims[1] : 0),
                 (unsigned long long)(rank > 2 ? gdims[2] : 0), (unsigned long long)(rank > 3 ? gdims[3] : 0), gbox[0],
                 rank > 1 ? gbox[1] : 0, rank > 2 ? gbox[2] : 0, rank > 3 ? gbox[3] : 0);
  return B200SD_OK;
}

int encode_tmap_act(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box, const uint32_t* elem_strides) {
  return encode_tmap_any(tm, kTmaDtype, base, rank, dims, strides_bytes, box, elem_strides);
}
int encode_tmap_f32(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box) {
  return encode_tmap_any(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_bytes, box, nullptr);
}

}  // namespace b200sd

using namespace b200sd;

extern "C" int b200sd_version(void) { return 100; }

extern "C" int b200sd_act_dtype(void) { return kActCode; }

extern "C" const char* b200sd_last_error(void) { return last_error_buf(); }

extern "C" int b200sd_device_info(int32_t* out4) {
  B200SD_REQUIRE(out4 != nullptr, B200SD_ERR_BAD_ARGUMENT, "device_info: out is NULL");
  DeviceInfo di;
  if (int e = get_device_info(&di)) return e;
  out4[0] = di.sm_count;
  out4[1] = di.cc_major;
  out4[2] = di.cc_minor;
  out4[3] = di.max_smem_optin;
  return B200SD_OK;
}

extern "C" int b200sd_struct_sizes(int32_t* out, int32_t max_entries) {
  const int32_t sizes[] = {
      (int32_t)sizeof(b200sd_gemm_params),       (int32_t)sizeof(b200sd_attn_params),
      (int32_t)sizeof(b200sd_groupnorm_params),  (int32_t)sizeof(b200sd_layernorm_params),
      (int32_t)sizeof(b200sd_prep_params),       (int32_t)sizeof(b200sd_temb_params),
      (int32_t)sizeof(b200sd_sched_params),      (int32_t)sizeof(b200sd_embed_params),
      (int32_t)sizeof(b200sd_denorm_params),     (int32_t)sizeof(b200sd_latent_cast_params),
      (int32_t)sizeof(b200sd_softmax_params),     (int32_t)sizeof(b200sd_memset_params),
  };
  const int n = (int)(sizeof(sizes) / sizeof(sizes[0]));
  for (int i = 0; i < n && i < max_entries; ++i) out[i] = sizes[i];
  return n;
}
