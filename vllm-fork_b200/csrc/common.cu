// b200ops -- error plumbing, device queries, TMA descriptor encoding.
#include "common.cuh"

#include <cstdlib>
#include <mutex>

namespace b200 {

static thread_local char g_err[1024] = {0};

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_last_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return B200_ERR_CUDA;
}

static int g_sm_count[64];
static int g_smem_optin[64];
static std::once_flag g_once;

static int cur_dev() {
  int d = -1;
  if (cudaGetDevice(&d) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return d;
}

int sm_count() {
  int d = cur_dev();
  if (d < 0 || d >= 64) return -1;
  if (g_sm_count[d] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, d) != cudaSuccess) return -1;
    g_sm_count[d] = v;
  }
  return g_sm_count[d];
}

int max_smem_optin() {
  int d = cur_dev();
  if (d < 0 || d >= 64) return -1;
  if (g_smem_optin[d] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, d) != cudaSuccess)
      return -1;
    g_smem_optin[d] = v;
  }
  return g_smem_optin[d];
}

// programmatic dependent launch level: 0 off, 1 launch overlap (default), 2 + weight prefetch before the wait.
// Initial value from B200_PDL; b200_set_pdl changes it.
static int g_pdl = -1;
int pdl_level() {
  if (g_pdl < 0) {
    const char* e = getenv("B200_PDL");
    g_pdl = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1;
  }
  return g_pdl;
}
bool pdl_enabled() { return pdl_level() > 0; }
int pdl_collective_mode() {
  static const int m = [] { const char* e = getenv("B200_PDL_AR"); return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2; }();
  return m;
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled g_encode = nullptr;

int encode_tensor_map(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base,
                      const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                      CUtensorMapSwizzle swz, CUtensorMapL2promotion l2) {
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
      set_last_error("cuTensorMapEncodeTiled not available (cudaGetDriverEntryPoint: %s)",
                     cudaGetErrorString(e));
      cudaGetLastError();
      return B200_ERR_CUDA;
    }
    g_encode = (PFN_encodeTiled)fn;
  }
  cuuint64_t gdims[5];
  cuuint64_t gstr[4];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; i++) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = g_encode(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdims, gstr, gbox, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error(
        "cuTensorMapEncodeTiled failed (CUresult %d): rank %d dims [%llu %llu %llu %llu %llu] "
        "strides [%llu %llu %llu %llu] box [%u %u %u %u %u]",
        (int)r, rank, (unsigned long long)gdims[0], (unsigned long long)(rank > 1 ? gdims[1] : 0),
        (unsigned long long)(rank > 2 ? gdims[2] : 0), (unsigned long long)(rank > 3 ? gdims[3] : 0),
        (unsigned long long)(rank > 4 ? gdims[4] : 0), (unsigned long long)(rank > 1 ? gstr[0] : 0),
        (unsigned long long)(rank > 2 ? gstr[1] : 0), (unsigned long long)(rank > 3 ? gstr[2] : 0),
        (unsigned long long)(rank > 4 ? gstr[3] : 0), gbox[0], rank > 1 ? gbox[1] : 0,
        rank > 2 ? gbox[2] : 0, rank > 3 ? gbox[3] : 0, rank > 4 ? gbox[4] : 0);
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // namespace b200

extern "C" {

const char* b200_last_error(void) { return b200::g_err; }

int b200_version(void) { return 1000; }

int b200_set_pdl(int level) {
  const int prev = b200::pdl_level();
  b200::g_pdl = level < 0 ? 0 : level > 2 ? 2 : level;
  return prev;
}

int b200_sm_count(void) {
  int n = b200::sm_count();
  if (n < 0) {
    b200::set_last_error("no CUDA device available (b200ops has no CPU fallback)");
    return B200_ERR_CUDA;
  }
  return n;
}

}  // extern "C"
