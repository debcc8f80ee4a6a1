// C-ABI housekeeping: error string, ABI version.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace scvi {
static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

// Per-kernel event profiler (scvi_profile_begin / _end): every launch of this library records an event
// behind itself on the profiled stream; a delay kernel in front keeps the queue full so consecutive
// events bracket one kernel's execution.  Not for use during stream capture.
struct ProfRec {
  std::string name;
  cudaEvent_t ev;
};
static bool g_prof = false;
static cudaStream_t g_prof_stream = nullptr;
static cudaEvent_t g_prof_start = nullptr;
static std::vector<ProfRec> g_prof_recs;

void count_launch(const char* name) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (g_prof) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) == cudaSuccess) {
      cudaEventRecord(e, g_prof_stream);
      g_prof_recs.push_back({name ? name : "?", e});
    }
  }
}

__global__ void prof_delay_kernel(long long ns) {
  long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 >= ns) break;
    __nanosleep(1000);
  }
}
int smem_optin(const void* func, int bytes, std::atomic<uint64_t>& done, const char* who) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  const uint64_t bit = 1ull << (dev & 63);
  if (done.load(std::memory_order_acquire) & bit) return SCVI_OK;
  cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) {
    set_error("%s: cannot reserve %d B of shared memory: %s", who, bytes, cudaGetErrorString(e));
    return SCVI_ECUDA;
  }
  done.fetch_or(bit, std::memory_order_release);
  return SCVI_OK;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

__global__ void __launch_bounds__(1024) colsum_kernel(const float* __restrict__ part, int64_t T, int64_t stride,
                                                     int64_t n, ColsumSegs segs) {
  __shared__ float red[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t e = (int64_t)blockIdx.x * 32 + tx;
  float s = 0.f;
  if (e < n) {
#pragma unroll 4
    for (int64_t t = ty; t < T; t += 32) s += part[t * stride + e];
  }
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && e < n) {
    float tot = 0.f;
#pragma unroll
    for (int g = 0; g < 32; ++g) tot += red[g][tx];
    for (int i = 0; i < segs.n; ++i)
      if (segs.dst[i] && e >= segs.begin[i] && e < segs.end[i]) segs.dst[i][e - segs.begin[i]] = tot;
  }
}

// few tiles, many columns: one thread per column
__global__ void colsum_flat_kernel(const float* __restrict__ part, int64_t T, int64_t stride, int64_t n,
                                   ColsumSegs segs) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float tot = 0.f;
#pragma unroll 4
  for (int64_t t = 0; t < T; ++t) tot += part[t * stride + e];
  for (int i = 0; i < segs.n; ++i)
    if (segs.dst[i] && e >= segs.begin[i] && e < segs.end[i]) segs.dst[i][e - segs.begin[i]] = tot;
}

int launch_colsum(const float* part, int64_t T, int64_t stride, int64_t n, const ColsumSegs& segs, cudaStream_t st,
                  const char* name) {
  if (n <= 0) return SCVI_OK;
  if (T <= 16 || n >= 65536) {
    colsum_flat_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(part, T, stride, n, segs);
    SCVI_LAUNCH_CHECK(name);
    return SCVI_OK;
  }
  colsum_kernel<<<(unsigned)cdiv(n, 32), 1024, 0, st>>>(part, T, stride, n, segs);
  SCVI_LAUNCH_CHECK(name);
  return SCVI_OK;
}
}  // namespace scvi

extern "C" {
int scvi_b200_abi_version(void) { return 3; }
const char* scvi_last_error_string(void) { return scvi::g_err; }
int64_t scvi_launch_count(void) { return (int64_t)scvi::g_launches.load(std::memory_order_relaxed); }

int scvi_host_device_pointer(const void* host_ptr, void** dev_ptr) {
  using namespace scvi;
  SCVI_REQUIRE(host_ptr && dev_ptr, "scvi_host_device_pointer: NULL pointer");
  void* d = nullptr;
  cudaError_t e = cudaHostGetDevicePointer(&d, const_cast<void*>(host_ptr), 0);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("scvi_host_device_pointer: %s (the buffer must be page-locked, mapped host memory)", cudaGetErrorString(e));
    return SCVI_ECUDA;
  }
  *dev_ptr = d;
  return SCVI_OK;
}

int scvi_profile_begin(void* stream, int64_t delay_us) {
  using namespace scvi;
  if (g_prof) {
    set_error("scvi_profile_begin: a profile is already open");
    return SCVI_EINVAL;
  }
  g_prof_stream = (cudaStream_t)stream;
  g_prof_recs.clear();
  if (delay_us > 0) prof_delay_kernel<<<1, 1, 0, g_prof_stream>>>(delay_us * 1000);
  if (cudaEventCreate(&g_prof_start) != cudaSuccess || cudaEventRecord(g_prof_start, g_prof_stream) != cudaSuccess) {
    set_error("scvi_profile_begin: %s", cudaGetErrorString(cudaGetLastError()));
    return SCVI_ECUDA;
  }
  g_prof = true;
  return SCVI_OK;
}

int64_t scvi_profile_end(char* buf, int64_t cap) {
  using namespace scvi;
  if (!g_prof) {
    set_error("scvi_profile_end: no open profile");
    return -1;
  }
  g_prof = false;
  std::vector<std::string> order;
  std::map<std::string, std::pair<long long, double>> agg;
  cudaEvent_t prev = g_prof_start;
  bool ok = true;
  if (!g_prof_recs.empty()) ok = cudaEventSynchronize(g_prof_recs.back().ev) == cudaSuccess;
  for (auto& r : g_prof_recs) {
    float ms = 0.f;
    if (ok && cudaEventElapsedTime(&ms, prev, r.ev) != cudaSuccess) ok = false;
    auto it = agg.find(r.name);
    if (it == agg.end()) {
      order.push_back(r.name);
      agg[r.name] = {1, (double)ms};
    } else {
      it->second.first += 1;
      it->second.second += ms;
    }
    if (prev != g_prof_start) cudaEventDestroy(prev);
    prev = r.ev;
  }
  if (prev != g_prof_start) cudaEventDestroy(prev);
  cudaEventDestroy(g_prof_start);
  g_prof_recs.clear();
  if (!ok) {
    cudaGetLastError();
    set_error("scvi_profile_end: event timing failed");
    return -1;
  }
  std::string out;
  char line[384];
  for (auto& n : order) {
    snprintf(line, sizeof(line), "%s\t%lld\t%.6f\n", n.c_str(), agg[n].first, agg[n].second);
    out += line;
  }
  if ((int64_t)out.size() + 1 > cap) {
    set_error("scvi_profile_end: buffer too small (%lld bytes needed)", (long long)out.size() + 1);
    return -1;
  }
  memcpy(buf, out.c_str(), out.size() + 1);
  return (int64_t)out.size();
}
}
