// sb_hostxfer.cpp -- host half of the grid transport of sb_copy_obs_to_host.
//
// OGM / DrivableAreaGridMap pixels are 0 or 255 (sensors.py:702-716, :744-758), so the grids
// cross PCIe as 1 bit per pixel (k_pack_bits in sb_api.cu) and are widened back to the uint8
// layout of StdObs (format_obs.py:108-110, :176-178) here, by a small pool of host threads that
// works on chunk c while chunk c + 1 is still in flight.  This is transport decoding only: no
// observation is computed on the CPU.
//
// Compiled by g++ (not nvcc) so that the AVX-512 body can use target attributes.
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <thread>
#include <pthread.h>
#include <sched.h>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace {

typedef void (*ExpandFn)(const uint8_t* bits, uint8_t* out, size_t n_bytes_in);

uint64_t g_lut[256];
std::once_flag g_lut_once;

void init_lut() {
  for (int v = 0; v < 256; ++v) {
    uint64_t w = 0;
    for (int b = 0; b < 8; ++b)
      if (v & (1 << b)) w |= (uint64_t)0xff << (8 * b);
    g_lut[v] = w;
  }
}

// portable body: one 8-byte store per input byte
void expand_lut(const uint8_t* bits, uint8_t* out, size_t n) {
  uint64_t* o = reinterpret_cast<uint64_t*>(out);
  for (size_t i = 0; i < n; ++i) o[i] = g_lut[bits[i]];
}

#if defined(__x86_64__)
// 64 pixels per instruction: mask -> byte vector, written with non-temporal stores (the output is
// not read again by this thread and is 8x the input, so it should not displace the cache)
__attribute__((target("avx512f,avx512bw"))) void expand_avx512(const uint8_t* bits, uint8_t* out, size_t n) {
  size_t i = 0;
  if ((reinterpret_cast<uintptr_t>(out) & 63) == 0) {
    for (; i + 8 <= n; i += 8) {
      uint64_t m;
      std::memcpy(&m, bits + i, 8);
      _mm512_stream_si512(reinterpret_cast<__m512i*>(out + i * 8), _mm512_movm_epi8((__mmask64)m));
    }
    _mm_sfence();
  } else {
    for (; i + 8 <= n; i += 8) {
      uint64_t m;
      std::memcpy(&m, bits + i, 8);
      _mm512_storeu_si512(out + i * 8, _mm512_movm_epi8((__mmask64)m));
    }
  }
  if (i < n) expand_lut(bits + i, out + i * 8, n - i);
}
#endif

// Incremental variants: `prev` holds the bit image this destination was last written from; only the
// 64-pixel words that differ are rewritten (and recorded in prev).  A road surface or an empty film that
// did not move writes nothing.
typedef void (*ExpandDiffFn)(const uint8_t* bits, uint8_t* prev, uint8_t* out, size_t n_bytes_in);

void expand_diff_lut(const uint8_t* bits, uint8_t* prev, uint8_t* out, size_t n) {
  uint64_t* o = reinterpret_cast<uint64_t*>(out);
  for (size_t i = 0; i < n; ++i)
    if (bits[i] != prev[i]) {
      o[i] = g_lut[bits[i]];
      prev[i] = bits[i];
    }
}

#if defined(__x86_64__)
__attribute__((target("avx512f,avx512bw"))) void expand_diff_avx512(const uint8_t* bits, uint8_t* prev, uint8_t* out,
                                                                     size_t n) {
  size_t i = 0;
  const bool aligned = (reinterpret_cast<uintptr_t>(out) & 63) == 0;
  // eight words (512 pixels) per compare: an unchanged stretch costs two loads and one predictable branch
  for (; i + 64 <= n; i += 64) {
    const __m512i mv = _mm512_loadu_si512(bits + i), pv = _mm512_loadu_si512(prev + i);
    unsigned ne = _mm512_cmpneq_epu64_mask(mv, pv);
    if (!ne) continue;
    _mm512_storeu_si512(prev + i, mv);
    alignas(64) uint64_t w[8];
    _mm512_store_si512(w, mv);
    while (ne) {
      const int k = __builtin_ctz(ne);
      ne &= ne - 1;
      const __m512i v = _mm512_movm_epi8((__mmask64)w[k]);
      if (aligned) _mm512_stream_si512(reinterpret_cast<__m512i*>(out + (i + 8 * (size_t)k) * 8), v);
      else _mm512_storeu_si512(out + (i + 8 * (size_t)k) * 8, v);
    }
  }
  for (; i + 8 <= n; i += 8) {
    uint64_t m, p;
    std::memcpy(&m, bits + i, 8);
    std::memcpy(&p, prev + i, 8);
    if (m == p) continue;
    const __m512i v = _mm512_movm_epi8((__mmask64)m);
    if (aligned) _mm512_stream_si512(reinterpret_cast<__m512i*>(out + i * 8), v);
    else _mm512_storeu_si512(out + i * 8, v);
    std::memcpy(prev + i, &m, 8);
  }
  if (aligned) _mm_sfence();
  if (i < n) expand_diff_lut(bits + i, prev + i, out + i * 8, n - i);
}
#endif

ExpandDiffFn pick_expand_diff() {
  std::call_once(g_lut_once, init_lut);
#if defined(__x86_64__)
  if (__builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512f")) return expand_diff_avx512;
#endif
  return expand_diff_lut;
}

ExpandFn pick_expand() {
  std::call_once(g_lut_once, init_lut);
#if defined(__x86_64__)
  if (__builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512f")) return expand_avx512;
#endif
  return expand_lut;
}

struct Piece {
  const uint8_t* bits;
  uint8_t* out;
  size_t n;
  uint8_t* prev;  // null: full write
};

}  // namespace

// A fixed pool; submit() queues pieces, wait() blocks until every queued piece is done.
class SbExpandPool {
 public:
  // first_cpu >= 0: worker i is pinned to the (first_cpu + i)-th CPU this process may run on (wrapping), so the
  // pools of the ranks of one box (one process per GPU) sit on disjoint cores instead of migrating over each other
  explicit SbExpandPool(int threads, int first_cpu = -1) : fn_(pick_expand()), diff_fn_(pick_expand_diff()) {
    if (threads < 1) threads = 1;
    std::vector<int> cpus;
    if (first_cpu >= 0) {
      cpu_set_t set;
      CPU_ZERO(&set);
      if (sched_getaffinity(0, sizeof(set), &set) == 0)
        for (int c = 0; c < CPU_SETSIZE; ++c)
          if (CPU_ISSET(c, &set)) cpus.push_back(c);
    }
    for (int i = 0; i < threads; ++i) {
      workers_.emplace_back([this] { run(); });
      if (!cpus.empty()) {
        cpu_set_t one;
        CPU_ZERO(&one);
        CPU_SET(cpus[(size_t)(first_cpu + i) % cpus.size()], &one);
        pthread_setaffinity_np(workers_.back().native_handle(), sizeof(one), &one);
      }
    }
  }
  ~SbExpandPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  void submit(const uint8_t* bits, uint8_t* out, size_t n, size_t piece, uint8_t* prev = nullptr) {
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t off = 0; off < n; off += piece) {
      queue_.push_back(Piece{bits + off, out + off * 8, n - off < piece ? n - off : piece, prev ? prev + off : nullptr});
      ++pending_;
    }
    cv_.notify_all();
  }
  void wait() {
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [this] { return pending_ == 0; });
  }
  int threads() const { return (int)workers_.size(); }

 private:
  void run() {
    for (;;) {
      Piece p;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [this] { return stop_ || head_ < queue_.size(); });
        if (stop_ && head_ >= queue_.size()) return;
        p = queue_[head_++];
      }
      if (p.prev) diff_fn_(p.bits, p.prev, p.out, p.n);
      else fn_(p.bits, p.out, p.n);
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (--pending_ == 0) {
          queue_.clear();
          head_ = 0;
          done_cv_.notify_all();
        }
      }
    }
  }
  ExpandFn fn_;
  ExpandDiffFn diff_fn_;
  std::vector<std::thread> workers_;
  std::vector<Piece> queue_;
  size_t head_ = 0, pending_ = 0;
  bool stop_ = false;
  std::mutex mu_;
  std::condition_variable cv_, done_cv_;
};

// C-linkage handles for sb_api.cu (which nvcc compiles)
#define SBX_LOCAL __attribute__((visibility("hidden")))
extern "C" {
SBX_LOCAL void* sbx_pool_create(int threads, int first_cpu);
SBX_LOCAL void sbx_pool_destroy(void* p);
SBX_LOCAL void sbx_pool_submit(void* p, const uint8_t* bits, uint8_t* out, size_t n_bytes_in, size_t piece);
SBX_LOCAL void sbx_pool_submit_diff(void* p, const uint8_t* bits, uint8_t* prev, uint8_t* out, size_t n_bytes_in,
                                    size_t piece);
SBX_LOCAL void sbx_pool_wait(void* p);
SBX_LOCAL int sbx_pool_threads(void* p);
void* sbx_pool_create(int threads, int first_cpu) { return new SbExpandPool(threads, first_cpu); }
void sbx_pool_destroy(void* p) { delete static_cast<SbExpandPool*>(p); }
void sbx_pool_submit(void* p, const uint8_t* bits, uint8_t* out, size_t n_bytes_in, size_t piece) {
  static_cast<SbExpandPool*>(p)->submit(bits, out, n_bytes_in, piece);
}
void sbx_pool_submit_diff(void* p, const uint8_t* bits, uint8_t* prev, uint8_t* out, size_t n_bytes_in, size_t piece) {
  static_cast<SbExpandPool*>(p)->submit(bits, out, n_bytes_in, piece, prev);
}
void sbx_pool_wait(void* p) { static_cast<SbExpandPool*>(p)->wait(); }
int sbx_pool_threads(void* p) { return static_cast<SbExpandPool*>(p)->threads(); }
}
