// Library-wide plumbing: last-error text, launch accounting, device properties.
#include "common.cuh"
#include <stdlib.h>

#include <stdarg.h>

#include <map>
#include <mutex>
#include <string>

namespace tadnn {
namespace {
thread_local char g_err[1024] = "";
std::mutex g_mu;
long long g_launches = 0;
std::map<std::string, long long>& table() {
  static std::map<std::string, long long> t;
  return t;
}
}  // namespace

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("TADNN_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

bool pdl_in_graphs() {
  static const bool on = [] {
    const char* e = getenv("TADNN_PDL_GRAPH");
    return !(e && e[0] == '0');
  }();
  return on;
}

int note_launch(const char* kernel_name) {
  std::lock_guard<std::mutex> lk(g_mu);
  ++g_launches;
  ++table()[kernel_name];
  return 0;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// Tiled bf16 tensor map of rank 2..5 with SWIZZLE_128B (box[0] must be 64 elements = 128 bytes).
// gdim / box are innermost-first; gstride_bytes has rank-1 entries (dimension 0 is contiguous).
int encode_tmap_bf16(CUtensorMap* out, const void* ptr, int rank, const unsigned long long* gdim,
                     const unsigned long long* gstride_bytes, const unsigned int* box) {
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  memset(out, 0, sizeof(*out));
  TADNN_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  TADNN_REQUIRE(rank >= 2 && rank <= 5, "encode_tmap_bf16: rank %d not in [2,5]", rank);
  TADNN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand %p is not 16-byte aligned", ptr);
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = gdim[i]; bx[i] = box[i]; es[i] = 1u; }
  for (int i = 0; i + 1 < rank; ++i) {
    TADNN_REQUIRE(gstride_bytes[i] % 16 == 0, "encode_tmap_bf16: stride %llu of dimension %d is not a multiple of 16 bytes", gstride_bytes[i], i + 1);
    gs[i] = gstride_bytes[i];
  }
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank), const_cast<void*>(ptr), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TADNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(rank %d) failed with CUresult %d", rank, static_cast<int>(r));
  return 0;
}

}  // namespace tadnn

extern "C" const char* tadnn_last_error(void) { return tadnn::g_err; }
extern "C" int tadnn_version(void) { return 100; }
extern "C" long long tadnn_launch_count(void) {
  std::lock_guard<std::mutex> lk(tadnn::g_mu);
  return tadnn::g_launches;
}
extern "C" void tadnn_reset_launch_count(void) {
  std::lock_guard<std::mutex> lk(tadnn::g_mu);
  tadnn::g_launches = 0;
  tadnn::table().clear();
}
extern "C" int tadnn_launch_table(char* host_buf, int cap) {
  std::lock_guard<std::mutex> lk(tadnn::g_mu);
  std::string s;
  for (auto& kv : tadnn::table()) s += kv.first + " " + std::to_string(kv.second) + "\n";
  if (cap <= 0 || host_buf == nullptr) return static_cast<int>(s.size());
  int n = static_cast<int>(s.size()) < cap - 1 ? static_cast<int>(s.size()) : cap - 1;
  memcpy(host_buf, s.data(), n);
  host_buf[n] = 0;
  return n;
}
