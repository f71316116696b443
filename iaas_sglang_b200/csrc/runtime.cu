// C-ABI plumbing: thread-local error string, device queries, TMA descriptor encoding.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("CUDA error in %s: %s", what, cudaGetErrorString(e));
  return B200_ERR_CUDA;
}

static unsigned long long g_launches = 0;

int launched(const char* what) {
  __atomic_add_fetch(&g_launches, 1ULL, __ATOMIC_RELAXED);
  return check_cuda(cudaGetLastError(), what);
}
unsigned long long launch_count() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

struct TuneEntry { const char* name; const char* env; int def; };
static const TuneEntry kTune[TUNE_COUNT] = {
    {"pdl", "B200_PDL", 1},               // programmatic dependent launch between the kernels of a step
    {"sk_max_m", "B200_SK_MAX_M", -1},    // FP8: >= 0 forces "skinny kernel iff M <= value"; -1 = measured (M, N, K) policy
    {"sk_stages", "B200_SK_STAGES", 0},   // cap on the skinny kernel's TMA ring depth (0 = as many as fit)
    {"sk_splits", "B200_SK_SPLITS", 0},   // force the split-K cluster size (0 = heuristic)
    {"sk_debug", "B200_SK_DEBUG", 0},
    {"gb_pair", "B200_GB_PAIR", 2},       // large-M kernel: 2 = cta_group::2 pairs where M allows
    {"gb_bn", "B200_GB_BN", 0},          // large-M fp8 tile width: 128 / 256, 0 = chosen by shape
    {"gb_prof", "B200_GB_PROF", 0},
    {"awq_nt", "B200_AWQ_NT", 0},
    {"awq_splits", "B200_AWQ_SPLITS", 0},
    {"awq_debug", "B200_AWQ_DEBUG", 0},
    {"awq_v1", "B200_AWQ_V1", 0},
    {"attn_ld_hint", "B200_ATTN_LD_HINT", 0},   // short-row KV loads: 0 = L2::256B prefetch hint, 1 = plain cp.async.cg
    {"gb_epi", "B200_GB_EPI", 0},
    {"ar_block_vec", "B200_AR_BLOCK_VEC", 0},   // push all-reduce: 16-byte vectors per block (0 = 256: one row per block); same on every rank
    {"ar_fused_world", "B200_AR_FUSED_WORLD", 4},   // largest world size at which all-reduce + add + RMSNorm runs as ONE kernel
    {"gb_split", "B200_GB_SPLIT", 0},   // large-M fp8 kernel, few-tile shapes: K slices per tile (0 = by shape, 1 = off)
};
static int g_tune[TUNE_COUNT];
static const bool g_tune_loaded = [] {
  for (int i = 0; i < TUNE_COUNT; ++i) {
    const char* e = getenv(kTune[i].env);
    g_tune[i] = e ? atoi(e) : kTune[i].def;
  }
  return true;
}();
int tune(TuneKey k) { return g_tune[k]; }
bool pdl_enabled() { return g_tune[TUNE_PDL] != 0; }

int device_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    cached[dev] = n;
  }
  return cached[dev];
}

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

int make_tmap_2d(CUtensorMap* map, const void* ptr, CUtensorMapDataType dt, int elem_bytes, uint64_t rows,
                 uint64_t cols, uint64_t ld_bytes, uint32_t box_rows, uint32_t box_cols, CUtensorMapSwizzle swz) {
  PFN_encodeTiled enc = get_encode_tiled();
  B200_REQUIRE(enc != nullptr, B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  B200_REQUIRE(((uintptr_t)ptr & 15) == 0 && ld_bytes % 16 == 0, B200_ERR_ARG,
               "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch (ptr=%p pitch=%llu)", ptr,
               (unsigned long long)ld_bytes);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  (void)elem_bytes;
  CUresult r = enc(map, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace b200

extern "C" const char* b200_last_error(void) { return b200::g_err; }
extern "C" int b200_abi_version(void) { return B200_SGL_ABI_VERSION; }

extern "C" int b200_sm_arch(void) {
  static int cached[64] = {0};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    b200::set_error("cudaGetDevice failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  if (dev >= 0 && dev < 64 && cached[dev]) return cached[dev];
  int major = 0, minor = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) {
    b200::set_error("cudaDeviceGetAttribute(compute capability) failed");
    return B200_ERR_CUDA;
  }
  const int arch = major * 10 + minor;
  if (dev >= 0 && dev < 64) cached[dev] = arch;
  return arch;
}

extern "C" int b200_tune_set(const char* name, int value) {
  for (int i = 0; i < b200::TUNE_COUNT; ++i)
    if (name && strcmp(name, b200::kTune[i].name) == 0) { b200::g_tune[i] = value; return B200_OK; }
  b200::set_error("b200_tune_set: unknown knob '%s'", name ? name : "(null)");
  return B200_ERR_ARG;
}
extern "C" int b200_tune_get(const char* name, int* value) {
  for (int i = 0; i < b200::TUNE_COUNT; ++i)
    if (name && value && strcmp(name, b200::kTune[i].name) == 0) { *value = b200::g_tune[i]; return B200_OK; }
  b200::set_error("b200_tune_get: unknown knob '%s'", name ? name : "(null)");
  return B200_ERR_ARG;
}

extern "C" int b200_sm_count(void) { return b200::device_sm_count(); }
namespace b200 { unsigned long long launch_count(); }
extern "C" unsigned long long b200_launch_count(void) { return b200::launch_count(); }
