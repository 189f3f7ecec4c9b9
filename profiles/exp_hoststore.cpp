// Host-memory write roofline for the grid widening of sb_copy_obs_to_host: N threads stream
// non-temporal 64-byte stores over a 2 GiB buffer (what the widened uint8 grids of one tick are).
//   g++ -O3 -pthread -mavx512f -mavx512bw profiles/exp_hoststore.cpp -o profiles/_bin/exp_hoststore
#include <immintrin.h>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
int main(int argc, char** argv) {
  int nt = argc > 1 ? atoi(argv[1]) : (int)std::thread::hardware_concurrency();
  size_t bytes = (size_t)2 << 30;
  uint8_t* buf = (uint8_t*)aligned_alloc(4096, bytes);
  for (size_t i = 0; i < bytes; i += 4096) buf[i] = 1;  // fault in
  for (int mode = 0; mode < 2; ++mode)
    for (int rep = 0; rep < 3; ++rep) {
      auto t0 = std::chrono::steady_clock::now();
      std::vector<std::thread> th;
      for (int t = 0; t < nt; ++t)
        th.emplace_back([=] {
          size_t lo = bytes / nt * t, hi = bytes / nt * (t + 1);
          __m512i v = _mm512_set1_epi8((char)(rep + 1));
          if (mode == 0) for (size_t i = lo; i < hi; i += 64) _mm512_stream_si512((__m512i*)(buf + i), v);
          else for (size_t i = lo; i < hi; i += 64) _mm512_store_si512((__m512i*)(buf + i), v);
          _mm_sfence();
        });
      for (auto& x : th) x.join();
      double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      printf("%s stores, %d threads: %.2f ms for 2 GiB = %.1f GB/s\n", mode == 0 ? "non-temporal" : "regular", nt, ms, bytes / ms / 1e6);
    }
  return buf[12345] == 77;
}
