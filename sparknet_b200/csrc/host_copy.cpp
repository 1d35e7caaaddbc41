// host_copy.cpp — host-side helpers for data producers (no CUDA): the multi-threaded copy / byte-to-float widening a
// data callback uses to fill the library's pinned staging buffer (include/ccaffe.h sparknet_host_copy,
// sparknet_host_widen).  Reference counterpart: the JVM callback of src/main/scala/libs/Net.scala:194-221 and
// src/main/java/libs/ByteImage.java:86-104, which widen and copy element by element.
//
// The widening writes four bytes for every byte it reads, into a buffer the DMA engine reads right afterwards and the
// CPU never again: the stores are non-temporal (no read-for-ownership of the destination lines, no cache pollution),
// which takes a third off the host memory traffic of the fp32 input path.
#include <immintrin.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

namespace {

__attribute__((target("avx2"))) void widen_avx2(float* d, const uint8_t* s, long n, bool is_signed) {
  long i = 0;
  // scalar head until the destination is 32-byte aligned
  for (; i < n && ((uintptr_t)(d + i) & 31); ++i) d[i] = is_signed ? (float)(int8_t)s[i] : (float)s[i];
  for (; i + 8 <= n; i += 8) {
    const __m128i b = _mm_loadl_epi64((const __m128i*)(s + i));
    const __m256i w = is_signed ? _mm256_cvtepi8_epi32(b) : _mm256_cvtepu8_epi32(b);
    _mm256_stream_ps(d + i, _mm256_cvtepi32_ps(w));
  }
  for (; i < n; ++i) d[i] = is_signed ? (float)(int8_t)s[i] : (float)s[i];
  _mm_sfence();
}
void widen_plain(float* d, const uint8_t* s, long n, bool is_signed) {
  if (is_signed) for (long i = 0; i < n; ++i) d[i] = (float)(int8_t)s[i];
  else for (long i = 0; i < n; ++i) d[i] = (float)s[i];
}

template <class F>
void parallel_chunks(long n, long granule, int threads, F work) {
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  const long chunk = ((n + threads - 1) / threads + granule - 1) / granule * granule;   // ceil: threads * chunk >= n
  if (threads == 1 || n < 4 * granule) { work(0, n); return; }
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) {
    const long lo = (long)t * chunk;
    if (lo >= n) break;
    pool.emplace_back(work, lo, std::min(n, lo + chunk));
  }
  work(0, std::min(n, chunk));
  for (auto& th : pool) th.join();
}

}  // namespace

extern "C" {
// memcpy on `threads` host threads (glibc's memcpy already streams large copies past the caches)
void sparknet_host_copy(void* dst, const void* src, long bytes, int threads) {
  if (bytes <= 0) return;
  if (bytes < (1 << 20)) { memcpy(dst, src, (size_t)bytes); return; }
  parallel_chunks(bytes, 4096, threads, [=](long lo, long hi) { memcpy((char*)dst + lo, (const char*)src + lo, (size_t)(hi - lo)); });
}
// n bytes -> n floats (signed Java bytes or 0..255) on `threads` host threads
void sparknet_host_widen(void* dst, const void* src, long n, int is_signed, int threads) {
  if (n <= 0) return;
  static const bool avx2 = __builtin_cpu_supports("avx2");
  auto work = [=](long lo, long hi) {
    if (avx2) widen_avx2((float*)dst + lo, (const uint8_t*)src + lo, hi - lo, is_signed != 0);
    else widen_plain((float*)dst + lo, (const uint8_t*)src + lo, hi - lo, is_signed != 0);
  };
  if (n < (1 << 18)) { work(0, n); return; }
  parallel_chunks(n, 1024, threads, work);
}
}  // extern "C"
