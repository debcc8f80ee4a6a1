// Host side of the host-streamed route: out[b, :] = X[idx[b], :] for a count matrix in (pinned) host memory, on a
// small persistent pool of host threads.  This is the reference's per-minibatch `data[indexes]` on the host
// (src/scvi/data/_anntorchdataset.py:143-152 under AnnDataLoader, src/scvi/dataloaders/_ann_dataloader.py:21-138),
// i.e. what feeds the H2D copy of a minibatch that does not live in HBM.  Rows are 1-4 KB and random: the copy is
// bound by DRAM latency, so every thread prefetches the rows a few iterations ahead.  No CUDA call in here.
#include <emmintrin.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace scvi {
namespace {

struct GatherJob {
  const uint8_t* X;
  int64_t ldx;  // bytes
  const int64_t* idx;
  int64_t B, row_bytes;
  uint8_t* out;
  int64_t ldo;  // bytes
};

class HostPool {
 public:
  explicit HostPool(int n) : stop_(false), gen_(0), pending_(0) {
    for (int i = 0; i < n; ++i) workers_.emplace_back([this, i] { run(i); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> l(m_);
      stop_ = true;
      ++gen_;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  int size() const { return (int)workers_.size(); }
  // rows [0, B) in chunks handed out by an atomic counter; the calling thread works too
  void gather(const GatherJob& j, int nthreads) {
    job_ = j;
    next_.store(0, std::memory_order_relaxed);
    const int helpers = std::max(0, std::min(nthreads - 1, size()));
    {
      std::lock_guard<std::mutex> l(m_);
      pending_ = helpers;
      active_ = helpers;
      ++gen_;
    }
    cv_.notify_all();
    work();
    std::unique_lock<std::mutex> l(m_);
    done_.wait(l, [this] { return pending_ == 0; });
  }

 private:
  // one row: 16-byte loads + non-temporal stores when both ends are 16-byte aligned (the destination is a pinned
  // staging buffer the copy engine reads next: no read-for-ownership, no cache pollution), memcpy otherwise
  static void copy_row(uint8_t* dst, const uint8_t* src, int64_t nbytes) {
    if ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0) {
      const int64_t nv = nbytes >> 4;
      const __m128i* s = reinterpret_cast<const __m128i*>(src);
      __m128i* d = reinterpret_cast<__m128i*>(dst);
      int64_t i = 0;
      for (; i + 4 <= nv; i += 4) {
        const __m128i a = _mm_load_si128(s + i), b = _mm_load_si128(s + i + 1), c = _mm_load_si128(s + i + 2),
                      e = _mm_load_si128(s + i + 3);
        _mm_stream_si128(d + i, a);
        _mm_stream_si128(d + i + 1, b);
        _mm_stream_si128(d + i + 2, c);
        _mm_stream_si128(d + i + 3, e);
      }
      for (; i < nv; ++i) _mm_stream_si128(d + i, _mm_load_si128(s + i));
      if (nbytes & 15) memcpy(dst + (nv << 4), src + (nv << 4), (size_t)(nbytes & 15));
    } else {
      memcpy(dst, src, (size_t)nbytes);
    }
  }
  void work() {
    const GatherJob& j = job_;
    constexpr int64_t CHUNK = 64;
    const int64_t AHEAD = ahead_;
    for (;;) {
      const int64_t b0 = next_.fetch_add(CHUNK, std::memory_order_relaxed);
      if (b0 >= j.B) break;
      const int64_t b1 = std::min(j.B, b0 + CHUNK);
      for (int64_t b = b0; b < std::min(b1, b0 + AHEAD); ++b) {
        const uint8_t* p = j.X + j.idx[b] * j.ldx;
        for (int64_t o = 0; o < j.row_bytes; o += 64) __builtin_prefetch(p + o, 0, 0);
      }
      for (int64_t b = b0; b < b1; ++b) {
        if (b + AHEAD < b1) {
          const uint8_t* p = j.X + j.idx[b + AHEAD] * j.ldx;
          for (int64_t o = 0; o < j.row_bytes; o += 64) __builtin_prefetch(p + o, 0, 0);
        }
        if (nt_) copy_row(j.out + b * j.ldo, j.X + j.idx[b] * j.ldx, j.row_bytes);
        else memcpy(j.out + b * j.ldo, j.X + j.idx[b] * j.ldx, (size_t)j.row_bytes);
      }
      _mm_sfence();
    }
  }
  void run(int id) {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> l(m_);
        cv_.wait(l, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
        if (id >= active_) continue;  // not part of this job
      }
      work();
      {
        std::lock_guard<std::mutex> l(m_);
        if (--pending_ == 0) done_.notify_one();
      }
    }
  }
  std::vector<std::thread> workers_;
  std::mutex m_;
  std::condition_variable cv_, done_;
  bool stop_;
  uint64_t gen_;
  int pending_, active_ = 0;
  GatherJob job_{};
 public:
  bool nt_ = true;     // measured: streaming stores help (0.75 vs 1.2 ms for 8192 x 2000 B on 8 threads) ...
  int64_t ahead_ = 0;  // ... software prefetch of the next rows does not (the hardware prefetchers cover a 2 KB row)
 private:
  std::atomic<int64_t> next_{0};
};

std::mutex g_pool_mutex;
HostPool* g_pool = nullptr;  // leaked on purpose: worker threads must not be joined from a static destructor

}  // namespace
}  // namespace scvi

extern "C" {

int scvi_host_gather_rows(const void* X, int64_t n_rows, int64_t row_bytes, int64_t ldx_bytes, const int64_t* idx,
                          int64_t B, void* out, int64_t ldo_bytes, int n_threads) {
  using namespace scvi;
  SCVI_REQUIRE(X && idx && out, "scvi_host_gather_rows: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && B >= 0 && row_bytes >= 0 && ldx_bytes >= row_bytes && ldo_bytes >= row_bytes,
               "scvi_host_gather_rows: bad shape");
  for (int64_t b = 0; b < B; ++b)
    SCVI_REQUIRE(idx[b] >= 0 && idx[b] < n_rows, "scvi_host_gather_rows: row index %lld out of range", (long long)idx[b]);
  if (B == 0 || row_bytes == 0) return SCVI_OK;
  const int hw = (int)std::max(1u, std::thread::hardware_concurrency());
  const int want = n_threads > 0 ? std::min(n_threads, 256) : hw;
  std::lock_guard<std::mutex> l(g_pool_mutex);  // one gather at a time per process
  if (!g_pool || g_pool->size() < want - 1) {
    HostPool* old = g_pool;
    g_pool = new HostPool(std::max(0, want - 1));
    delete old;
  }
  GatherJob j{(const uint8_t*)X, ldx_bytes, idx, B, row_bytes, (uint8_t*)out, ldo_bytes};
  if (const char* e = getenv("SCVI_HOST_GATHER")) {  // experiment knobs: "<nt 0/1>,<rows ahead>"
    g_pool->nt_ = e[0] == '1';
    if (e[1] == ',') g_pool->ahead_ = atoi(e + 2);
  }
  g_pool->gather(j, want);
  return SCVI_OK;
}
}
