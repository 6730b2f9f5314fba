// Host-side expansion of the bit rows of dmg_neighbors_batch_bits into CSR-shaped (id, hop) rows.
// Pure host code (no CUDA): a consumer that wants arrays pays for writing ~5 bytes per neighbour here
// instead of moving them over PCIe; a consumer that walks the neighbours can read the planes directly.
#include <immintrin.h>

#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <thread>
#include <vector>

#include "../../include/dmgraph.h"

namespace dmg {
void set_error(const char* fmt, ...);
}

namespace {

// One row, scalar: every set bit of the member word becomes (id, code).
void decode_row_scalar(uint32_t planes, uint32_t plane_words, uint32_t lo, uint32_t nw, const uint32_t* row,
                       uint32_t* ids, uint8_t* hop) {
  for (uint32_t j = 0; j < nw; ++j) {
    uint32_t m = 0;
    for (uint32_t p = 0; p < planes; ++p) m |= row[(size_t)p * plane_words + j];
    while (m) {
      const int i = __builtin_ctz(m);
      m &= m - 1;
      uint32_t code = 0;
      for (uint32_t p = 0; p < planes; ++p) code |= ((row[(size_t)p * plane_words + j] >> i) & 1u) << p;
      *ids++ = lo + 32 * j + (uint32_t)i;
      *hop++ = (uint8_t)code;
    }
  }
}

// One row with AVX-512 compress stores: 32 vertices per step -- the member word is the mask of two
// 16-lane id compresses and of one 32-byte code compress (codes = plane bits expanded to bytes).
__attribute__((target("avx512f,avx512bw,avx512vl,avx512vbmi2"))) void decode_row_avx512(
    uint32_t planes, uint32_t plane_words, uint32_t lo, uint32_t nw, const uint32_t* row, uint32_t* ids, uint8_t* hop) {
  const __m512i iota = _mm512_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
  for (uint32_t j = 0; j < nw; ++j) {
    uint32_t m = 0;
    __m256i code = _mm256_setzero_si256();
    for (uint32_t p = 0; p < planes; ++p) {
      const uint32_t w = row[(size_t)p * plane_words + j];
      m |= w;
      code = _mm256_or_si256(code, _mm256_maskz_set1_epi8((__mmask32)w, (char)(1u << p)));
    }
    if (!m) continue;
    const __m512i base = _mm512_add_epi32(iota, _mm512_set1_epi32((int)(lo + 32 * j)));
    const __mmask16 m0 = (__mmask16)(m & 0xFFFFu), m1 = (__mmask16)(m >> 16);
    _mm512_mask_compressstoreu_epi32(ids, m0, base);
    ids += __builtin_popcount(m0);
    _mm512_mask_compressstoreu_epi32(ids, m1, _mm512_add_epi32(base, _mm512_set1_epi32(16)));
    ids += __builtin_popcount(m1);
    _mm256_mask_compressstoreu_epi8(hop, (__mmask32)m, code);
    hop += __builtin_popcount(m);
  }
}

// Streaming form for large results: a thread expands its rows into a small staging buffer that stays in
// its cache and moves it to the destination with non-temporal 64-byte stores -- the destination lines are
// written once and never read, so fetching them for ownership (which a normal store does) only doubles
// the memory traffic of a job that is bound by exactly that.  `flush` keeps the destination cursor on a
// 64-byte boundary between calls: what does not fill a line waits in the staging buffer for the next one.
__attribute__((target("clflushopt"))) inline void evict_lines(const void* p, size_t bytes) {
  const char* c = reinterpret_cast<const char*>(reinterpret_cast<uintptr_t>(p) & ~uintptr_t(63));
  const char* e = reinterpret_cast<const char*>(p) + bytes;
  for (; c < e; c += 64) _mm_clflushopt(const_cast<char*>(c));
}

template <class T>
struct NtStream {
  static constexpr size_t LINE = 64 / sizeof(T);
  T* dst;            // next destination element
  T* stage;          // staging buffer (capacity given by the caller)
  size_t n = 0;      // staged elements
  __attribute__((target("avx512f"))) void flush(bool final) {
    size_t i = 0;
    while (i < n && (reinterpret_cast<uintptr_t>(dst) & 63)) *dst++ = stage[i++];   // up to the first line boundary
    for (; i + LINE <= n; i += LINE, dst += LINE)
      _mm512_stream_si512(reinterpret_cast<__m512i*>(dst), _mm512_loadu_si512(stage + i));
    if (final) {
      while (i < n) *dst++ = stage[i++];
      n = 0;
    } else {
      const size_t rest = n - i;
      for (size_t k = 0; k < rest; ++k) stage[k] = stage[i + k];
      n = rest;
    }
  }
};

}  // namespace

namespace dmg {

// Rows [q_begin, q_end) of a bit-row result into out_ids / out_hop at out_off[q] (out_off already holds the
// offsets of these rows and of row q_end).  All arrays are indexed by absolute query.  Used by dmg_bits_decode
// and by dmg_neighbors_batch, which expands the chunks of a batch while later chunks are still on the device.
void bits_decode_rows(uint32_t planes, uint32_t plane_words, uint64_t q_begin, uint64_t q_end, const uint32_t* win,
                      const uint32_t* counts, const uint32_t* bits, const uint64_t* out_off, uint32_t* out_ids,
                      uint8_t* out_hop, int32_t n_threads) {
  if (q_end <= q_begin) return;
  const uint64_t nq = q_end - q_begin, total = out_off[q_end] - out_off[q_begin];
  int T = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  T = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)std::max(T, 1), nq / 64 + 1));
  const size_t row_words = (size_t)planes * plane_words;
  static const bool simd = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                           __builtin_cpu_supports("avx512vl") && __builtin_cpu_supports("avx512vbmi2");
  static const bool streaming = [] {
    const char* e = getenv("DMG_DECODE_NT");   // "0": plain stores (A/B runs)
    return !(e && e[0] == '0');
  }();
  // A consumed bit row is flushed from the caches (clflushopt): the pinned buffer it lies in is the target of
  // the caller's next device-to-host copy, and that copy was measured 3x slower (3.0 vs 0.96 ms for 34 MB on
  // the B200 box) when its lines still sat in the cores' caches.  DMG_DECODE_EVICT=0 turns it off (A/B runs).
  static const bool evict = [] {
    const char* e = getenv("DMG_DECODE_EVICT");
    return !(e && e[0] == '0') && __builtin_cpu_supports("clflushopt");
  }();
  constexpr size_t STAGE = 8192;   // entries staged per thread (32 KB of ids + 8 KB of hops: stays in L2)
  auto work_nt = [&](uint64_t a, uint64_t b) {
    std::vector<uint32_t> sid(STAGE + 64);
    std::vector<uint8_t> shp(STAGE + 64);
    NtStream<uint32_t> si{out_ids + out_off[a], sid.data()};
    NtStream<uint8_t> sh{out_hop + out_off[a], shp.data()};
    for (uint64_t q = a; q < b; ++q) {
      const uint32_t cnt = counts[q];
      if (!cnt) continue;
      if (si.n + cnt > STAGE || cnt > STAGE) {
        si.flush(false);
        sh.flush(false);
      }
      const uint32_t lo = win[2 * q], nw = (win[2 * q + 1] + 31) >> 5;
      const uint32_t* row = bits + q * row_words;
      if (cnt > STAGE - 64) {   // a row larger than the staging buffer goes straight to its place
        si.flush(true);
        sh.flush(true);
        decode_row_avx512(planes, plane_words, lo, nw, row, si.dst, sh.dst);
        si.dst += cnt;
        sh.dst += cnt;
        continue;
      }
      decode_row_avx512(planes, plane_words, lo, nw, row, si.stage + si.n, sh.stage + sh.n);
      si.n += cnt;
      sh.n += cnt;
      if (evict) evict_lines(row, row_words * 4);
    }
    si.flush(true);
    sh.flush(true);
    _mm_sfence();
  };
  auto work = [&](uint64_t a, uint64_t b) {
    if (simd && streaming && out_off[b] - out_off[a] >= (1u << 16)) return work_nt(a, b);
    for (uint64_t q = a; q < b; ++q) {
      if (!counts[q]) continue;
      const uint32_t lo = win[2 * q], nw = (win[2 * q + 1] + 31) >> 5;
      const uint32_t* row = bits + q * row_words;
      uint32_t* ids = out_ids + out_off[q];
      uint8_t* hop = out_hop + out_off[q];
      if (simd) decode_row_avx512(planes, plane_words, lo, nw, row, ids, hop);
      else decode_row_scalar(planes, plane_words, lo, nw, row, ids, hop);
    }
  };
  if (T == 1) {
    work(q_begin, q_end);
  } else {
    // split by output volume, not by query count
    std::vector<std::thread> th;
    uint64_t a = q_begin;
    for (int t = 0; t < T; ++t) {
      const uint64_t want = out_off[q_begin] + total / T * (t + 1);
      uint64_t b = t == T - 1 ? q_end
                              : (uint64_t)(std::upper_bound(out_off + q_begin, out_off + q_end + 1, want) - out_off);
      b = std::min(std::max(b, a), q_end);
      if (b > a) th.emplace_back(work, a, b);
      a = b;
    }
    for (auto& x : th) x.join();
  }
}

}  // namespace dmg

extern "C" int32_t dmg_bits_decode(uint32_t planes, uint32_t plane_words, uint64_t nq, const uint32_t* win,
                                   const uint32_t* counts, const uint32_t* bits, uint64_t* out_off, uint32_t* out_ids,
                                   uint8_t* out_hop, uint64_t cap, uint64_t* needed, int32_t n_threads) {
  if (!out_off || !needed || (nq && (!win || !counts || !bits)) || planes < 1 || planes > 8) {
    dmg::set_error("dmg_bits_decode: bad argument");
    return DMG_EINVAL;
  }
  uint64_t total = 0;
  for (uint64_t q = 0; q < nq; ++q) {
    out_off[q] = total;
    total += counts[q];
  }
  out_off[nq] = total;
  *needed = total;
  if (total > cap || (total && (!out_ids || !out_hop))) {
    dmg::set_error("dmg_bits_decode: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    return DMG_ECAP;
  }
  dmg::bits_decode_rows(planes, plane_words, 0, nq, win, counts, bits, out_off, out_ids, out_hop, n_threads);
  return DMG_OK;
}
