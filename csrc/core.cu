// Library-wide plumbing: last-error text, launch accounting, device properties.
#include "common.cuh"

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
