// S1 at scale (BASELINE config 3): every query of a batch is matched in ONE pass
// over the packed, pre-lowered name arena.
//
//   - the batch's distinct patterns are classed by length and keyed by their first 8 / 4 / 3 /
//     2 / 1 bytes; per class a Bloom-style bitmap in shared memory rejects almost every text
//     position, a small open-addressing table in L2 maps a surviving key to its patterns;
//   - one warp owns a contiguous tile of names; its 32 lanes take 32 consecutive names, ONE LANE PER
//     NAME, each sliding a 20-byte window over its own name (aligned 4-byte loads + funnel shifts), so
//     the arena is read once whatever the number of queries.  A position that passes a filter is only
//     queued; the warp resolves the queued candidates together afterwards (see scan_candidates /
//     resolve_candidates below);
//   - ordered output without a sort: pass 1 counts hits per (tile, pattern); a column scan turns
//     the counts into per-tile row cursors; pass 2 repeats the match and writes each hit at its
//     cursor.  Tiles are name ranges in id order and a warp walks its tile in id order, so every
//     row comes out in ascending id == `nodes` order (CCS:1511).  A pattern that occurs twice in
//     one name is reported once (String.contains);
//   - the "first id containing the query" scan of resolveClassName (GQS:449-453) is the same
//     match with an atomicMin per pattern instead of rows.
#include <algorithm>
#include <string>
#include <unordered_map>

#include "common.cuh"

namespace dmg {

constexpr int SM_THREADS = 256;          // 8 warps per CTA
constexpr int SM_WARPS = SM_THREADS / 32;
constexpr int SM_HITBUF = 32;            // distinct patterns of one name before the exhaustive fallback
constexpr uint32_t SM_MAX_PATTERNS = 16384;   // per round (bounds the tile x pattern matrix)

constexpr uint32_t PAT_NIL = 0xFFFFFFFFu;
struct __align__(16) PatSlot {     // one probe = one 16-byte load
  unsigned long long key;
  uint32_t first;                  // first pattern with this key, PAT_NIL = empty slot
  uint32_t fp;                     // fingerprint the key was inserted with (compared too)
};
struct __align__(16) PatInfo {     // one 16-byte load per candidate pattern
  unsigned long long pk8;          // first min(len, 8) bytes, little endian
  uint32_t len;
  uint32_t next;                   // next pattern with the same key, PAT_NIL = end of chain
};
struct __align__(16) PatRow {      // where a pattern's hits go (single-query patterns: one load)
  unsigned long long row_base;     // out_off of the pattern's first query
  uint32_t n_queries;
  uint32_t uq_begin;
};
struct SlotTable {
  const PatSlot* slot;             // [size]
  uint32_t mask;                   // size - 1
  uint32_t fbits;                  // log2 of the filter size in bits
  uint32_t present;
};

struct SearchArgs {
  const unsigned long long* name_off;
  const uint8_t* lower;
  uint32_t v_lo, v_hi, per_tile, n_tiles;
  SlotTable t16, t12, t8, t4, t3;   // classes by pattern length: >=16, 12..15, 8..11, 4..7, 3
  const uint32_t* filt;             // filters of the five classes, then the shared 4-byte prefix filter (words)
  uint32_t filt_words16, filt_words12, filt_words8, filt_words4, filt_words3, filt_wordsP;
  uint32_t fbitsP, presentP;        // prefix filter: 4-byte prefixes of every pattern of >= 4 bytes
  unsigned long long lower_size;    // bytes of the arena that may be read (names + zero pad)
  const PatInfo* pat;               // [U]
  PatRow* prow;                     // [U] filled by search_rows_kernel once out_off is known
  const uint32_t* poff;
  const uint8_t* pbytes;
  const uint32_t* tiny2;            // [65536] pattern id + 1
  const uint32_t* tiny1;            // [256]
  uint32_t has2, has1, empty_plus1;
  uint32_t U;                       // patterns in this round
  uint32_t* cnt;                    // [U][n_tiles] counts, then cursors (a pattern's tiles are contiguous)
  uint32_t* first;                  // [U] min id (MODE 2)
  const uint32_t* uq_off;           // pattern -> queries that asked for it
  const uint32_t* uq;
  const unsigned long long* out_off;
  uint32_t* out_ids;
  unsigned long long cap;
  // hit log of the count pass (see search_multi_kernel)
  uint32_t* log;
  unsigned long long log_cap;
  unsigned long long* log_used;
  unsigned long long* step_base;
  uint32_t* step_cnt;
  uint32_t steps_per_tile;
  uint32_t* tile_flag;              // [n_tiles] 1 = some step of the tile is not in the log (emit re-matches the tile)
};

// A log entry carries everything the emit pass needs to place the hit without looking at its neighbours:
// pattern (14 bits: SM_MAX_PATTERNS), name inside the 32-name step (5 bits) and the hit's rank among the
// tile's hits of that pattern (13 bits: a tile holds at most LOG_MAX_TILE names) -- the value the count
// pass's own counter returned.
constexpr uint32_t LOG_U_BITS = 14, LOG_LANE_BITS = 5, LOG_IDX_SHIFT = LOG_U_BITS + LOG_LANE_BITS;
constexpr uint32_t LOG_MAX_TILE = 1u << (32 - LOG_IDX_SHIFT);
static_assert(SM_MAX_PATTERNS <= (1u << LOG_U_BITS), "pattern ids must fit the log entry");
__device__ __forceinline__ uint32_t log_pack(uint32_t idx, uint32_t lane, uint32_t u) {
  return (idx << LOG_IDX_SHIFT) | (lane << LOG_U_BITS) | u;
}

// Everything on the per-position path is 32-bit arithmetic (64-bit multiplies cost several
// instructions each).  A class turns its window words into a table key (64 bits, compared exactly)
// and a 32-bit fingerprint; the filter index and the slot index are cheap functions of the fingerprint.
struct KeyFp {
  unsigned long long key;
  uint32_t fp;    // compared in the slot table
  uint32_t raw;   // what the filters are indexed by
};
// The per-position cost is what bounds the scan (ncu, config 3: ~185 thread instructions per byte
// position before, most of them the five classes' hashes), so the classes share one incremental
// hash: raw(4 bytes) = lo * A, raw(8) = raw(4) ^ hi * B, raw(12) = raw(8) ^ lo2 * C, raw(16) = raw(12) ^
// hi2 * D -- one multiply per class -- and the first filter bit is simply raw's top bits.  Key and
// fingerprint are only formed for the few positions that pass a filter.
constexpr uint32_t HASH_A = 0x9E3779B1u, HASH_B = 0x85EBCA77u, HASH_C = 0xC2B2AE3Du, HASH_D = 0x27D4EB2Fu;
__host__ __device__ __forceinline__ uint32_t class_raw(int cls, uint32_t lo, uint32_t hi, uint32_t lo2, uint32_t hi2) {
  switch (cls) {
    case 0: return (lo * HASH_A) ^ (hi * HASH_B) ^ (lo2 * HASH_C) ^ (hi2 * HASH_D);   // >= 16 bytes: first 16
    case 1: return (lo * HASH_A) ^ (hi * HASH_B) ^ (lo2 * HASH_C);                    // 12..15: first 12
    case 2: return (lo * HASH_A) ^ (hi * HASH_B);                                     // 8..11: first 8
    case 3: return lo * HASH_A;                                                       // 4..7: first 4
    default: return (lo & 0xFFFFFFu) * HASH_A;                                        // exactly 3
  }
}
__host__ __device__ __forceinline__ KeyFp class_key(int cls, uint32_t lo, uint32_t hi, uint32_t lo2, uint32_t hi2) {
  KeyFp r;
  switch (cls) {
    case 0: r.key = ((unsigned long long)(hi ^ (lo2 * 0x85EBCA6Bu)) << 32) | (lo ^ (hi2 * 0xC2B2AE35u)); break;
    case 1: r.key = ((unsigned long long)(hi ^ (lo2 * 0x85EBCA6Bu)) << 32) | lo; break;
    case 2: r.key = ((unsigned long long)hi << 32) | lo; break;
    case 3: r.key = lo; break;
    default: r.key = lo & 0xFFFFFFu;
  }
  r.raw = class_raw(cls, lo, hi, lo2, hi2);
  r.fp = r.raw ^ (r.raw >> 15);
  return r;
}
__host__ __device__ __forceinline__ uint32_t hash_slot(uint32_t fp, uint32_t mask) {
  return ((fp * 0x2C1B3C6Du) >> 9) & mask;
}
__host__ __device__ __forceinline__ uint32_t hash_filter(uint32_t raw, uint32_t bits) { return raw >> (32 - bits); }
// second, independent bit of the same filter: tested only when the first is set, so it costs nothing on
// the 9 positions in 10 the first bit rejects, and squares the false-positive rate (ncu, config 3: a false
// candidate pays a dependent 16-byte table load; at 10 filter bits per key that was ~18 per name)
__host__ __device__ __forceinline__ uint32_t hash_filter2(uint32_t raw, uint32_t bits) {
  return ((raw ^ (raw >> 13)) * 0x7FEB352Du) >> (32 - bits);
}
__device__ __forceinline__ bool filter_pass(const uint32_t* f, uint32_t raw, uint32_t bits) {
  const uint32_t fi = hash_filter(raw, bits);
  if (!((f[fi >> 5] >> (fi & 31)) & 1u)) return false;
  const uint32_t fj = hash_filter2(raw, bits);
  return (f[fj >> 5] >> (fj & 31)) & 1u;
}
__device__ __forceinline__ uint32_t table_find(const SlotTable& t, unsigned long long k, uint32_t fp) {
  uint32_t h = hash_slot(fp, t.mask);
  while (true) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(t.slot + h));
    if (raw.z == PAT_NIL) return PAT_NIL;
    if ((((unsigned long long)raw.y << 32) | raw.x) == k && raw.w == fp) return raw.z;
    h = (h + 1) & t.mask;
  }
}
__device__ __forceinline__ PatInfo load_pat(const PatInfo* p, uint32_t u) {
  const uint4 raw = __ldg(reinterpret_cast<const uint4*>(p + u));
  PatInfo r;
  r.pk8 = ((unsigned long long)raw.y << 32) | raw.x;
  r.len = raw.z;
  r.next = raw.w;
  return r;
}
__device__ __forceinline__ unsigned long long low_bytes_mask(uint32_t len) {
  return len >= 8 ? ~0ull : ((1ull << (8 * len)) - 1ull);
}

// Matching is two phases per warp step of 32 consecutive names, because the hit path -- slot table
// probe, pattern record, verification -- diverges hopelessly when it runs inline with the scan:
//   scan     the warp sweeps the step's CONTIGUOUS byte range (names are packed back to back), 128
//            bytes per iteration: a lane loads one aligned word (coalesced), gets the next four from its
//            neighbours by shuffle and tests the four byte positions of its word.  All 32 lanes work
//            whatever the lengths of the names (one lane per name left half of them idle: ncu counted
//            16.5 active threads per instruction).  A position first asks ONE shared filter keyed on
//            its 4 bytes (the 4-byte prefixes of every pattern of >= 4 bytes) and the 3-byte filter,
//            both in shared memory; only a position that passes goes on to the length classes' own
//            filters (global memory, L1-resident).  A survivor looks up the name it lies in (bisection
//            over the step's 33 offsets in shared memory) and is QUEUED as a candidate (class, position)
//            of that name if the name still has room for the class's shortest pattern there
//   resolve  the warp flattens the 32 queues and works them off 32 candidates at a time, all lanes
//            busy: window bytes -> key -> slot table -> chain of pattern records -> verification; a
//            match is appended to the hit list of the name it belongs to
// after which every lane de-duplicates its own short list (a pattern may occur at several positions).
constexpr int SM_CANDBUF = 40;                 // queued candidates of one name before the exhaustive fallback
constexpr uint32_t CAND_POS_BITS = 13;         // candidate = class << 13 | position; longer names: fallback

__device__ __forceinline__ bool filter_pass_ldg(const uint32_t* __restrict__ f, uint32_t raw, uint32_t bits) {
  const uint32_t fi = hash_filter(raw, bits);
  if (!((__ldg(&f[fi >> 5]) >> (fi & 31)) & 1u)) return false;
  const uint32_t fj = hash_filter2(raw, bits);
  return (__ldg(&f[fj >> 5]) >> (fj & 31)) & 1u;
}

// Both probes of a filter without a branch between them (inside the sweep a branch would only make the
// warp run both sides: some lane nearly always passes the first bit).
__device__ __forceinline__ bool filter_pass2(const uint32_t* f, uint32_t raw, uint32_t bits) {
  const uint32_t fi = hash_filter(raw, bits), fj = hash_filter2(raw, bits);
  return ((f[fi >> 5] >> (fi & 31)) & (f[fj >> 5] >> (fj & 31)) & 1u) != 0;
}

constexpr uint32_t SM_WORK = 192;          // positions a warp parks before it works them off
constexpr uint32_t WORK_P4 = 1u, WORK_3 = 2u, WORK_2 = 4u, WORK_1 = 8u;

// phase 1b, whole warp: the parked positions, 32 at a time with every lane busy.  A position that passed
// the prefix filter asks the length classes' own filters (global memory, L1-resident); what survives
// finds its name by bisection over the step's offsets and is queued there.
__device__ __forceinline__ void drain_work(const SearchArgs& a, int lane, unsigned long long B0, const uint32_t* work,
                                           uint32_t n_work, const uint32_t* s_start, uint32_t* s_nc,
                                           uint16_t (*cand)[SM_CANDBUF]) {
  const uint32_t* __restrict__ g16 = a.filt;
  const uint32_t* __restrict__ g12 = g16 + a.filt_words16;
  const uint32_t* __restrict__ g8 = g12 + a.filt_words12;
  const uint32_t* __restrict__ g4 = g8 + a.filt_words8;
  for (uint32_t i = lane; i < n_work; i += 32) {
    const uint32_t item = work[i], r0 = item >> 4, f = item & 15u;
    uint32_t cm = (f & WORK_3 ? 1u << 4 : 0u) | (f & WORK_2 ? 1u << 5 : 0u) | (f & WORK_1 ? 1u << 6 : 0u);
    if (f & WORK_P4) {
      const unsigned long long at = B0 + r0;
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(a.lower + (at & ~3ull));   // zero padded past the end
      const uint32_t sh = (uint32_t)(at & 3ull) * 8;
      const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3), w4 = __ldg(wp + 4);
      const uint32_t lo32 = sh ? __funnelshift_r(w0, w1, sh) : w0, hi32 = sh ? __funnelshift_r(w1, w2, sh) : w1;
      const uint32_t lo2 = sh ? __funnelshift_r(w2, w3, sh) : w2, hi2 = sh ? __funnelshift_r(w3, w4, sh) : w3;
      const uint32_t h4 = lo32 * HASH_A, h8 = h4 ^ (hi32 * HASH_B), h12 = h8 ^ (lo2 * HASH_C);
      if (a.t4.present && filter_pass_ldg(g4, h4, a.t4.fbits)) cm |= 1u << 3;
      if (a.t8.present && filter_pass_ldg(g8, h8, a.t8.fbits)) cm |= 1u << 2;
      if (a.t12.present && filter_pass_ldg(g12, h12, a.t12.fbits)) cm |= 1u << 1;
      if (a.t16.present && filter_pass_ldg(g16, h12 ^ (hi2 * HASH_D), a.t16.fbits)) cm |= 1u;
    }
    if (!cm) continue;
    uint32_t l = 0;   // the name the position lies in
#pragma unroll
    for (int st = 16; st > 0; st >>= 1)
      if (s_start[l + st] <= r0) l += st;
    const uint32_t pos = r0 - s_start[l], room = s_start[l + 1] - r0;
    for (; cm; cm &= cm - 1) {
      const uint32_t cls = __ffs(cm) - 1;
      const uint32_t shortest = cls == 0 ? 16u : cls == 1 ? 12u : cls == 2 ? 8u : cls == 3 ? 4u : 7u - cls;   // 3, 2, 1
      if (room < shortest) continue;
      const uint32_t slot = atomicAdd(&s_nc[l], 1u);
      if (slot < (uint32_t)SM_CANDBUF) cand[l][slot] = (uint16_t)((cls << CAND_POS_BITS) | pos);
    }
  }
  __syncwarp();
}

// phase 1, whole warp.  B0 / B1 = byte range of the step's names; s_start[l] = start of name l relative to
// B0 (entries past the last name and s_start[32] hold B1 - B0); s_nc[l] = queue length of name l (may run
// past SM_CANDBUF: overflow; preset beyond it for names that cannot be queued at all).  The sweep itself
// only asks the two shared-memory filters and PARKS the few positions that pass (work[], SM_WORK entries):
// anything rarer than that would run with a handful of lanes if it ran inline.
template <bool TINY>   // TINY: the batch has patterns of one or two bytes (their tables are asked per position)
__device__ __forceinline__ void scan_step(const SearchArgs& a, const uint32_t* fP, const uint32_t* f3, int lane,
                                          unsigned long long B0, unsigned long long B1, const uint32_t* s_start,
                                          uint32_t* s_nc, uint16_t (*cand)[SM_CANDBUF], uint32_t* work) {
  const unsigned long long c0 = B0 & ~3ull;
  const uint32_t* __restrict__ wbase = reinterpret_cast<const uint32_t*>(a.lower + c0);
  const unsigned long long n_words = (a.lower_size - c0) >> 2;   // words that may be read from c0 on
  const uint32_t span = (uint32_t)(B1 - B0);
  const unsigned lt = lanemask_lt();
  uint32_t nxt = lane < n_words ? __ldg(wbase + lane) : 0u;
  uint32_t n_work = 0;
  unsigned long long wi = 0;
  for (unsigned long long c = c0; c < B1; c += 128, wi += 32) {
    const uint32_t w0 = nxt;
    nxt = wi + 32 + lane < n_words ? __ldg(wbase + wi + 32 + lane) : 0u;
    // the word after the lane's own: the neighbour's, or for lane 31 the first word of the next chunk
    const uint32_t w1 = __shfl_sync(FULL, lane >= 1 ? w0 : nxt, (lane + 1) & 31);
    const int rel0 = (int)((long long)(c + 4ull * lane) - (long long)B0);   // first position of the lane, relative to B0
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int rel = rel0 + j;
      const uint32_t lo32 = j ? __funnelshift_r(w0, w1, 8 * j) : w0;
      uint32_t f = 0;
      if (a.presentP && filter_pass2(fP, lo32 * HASH_A, a.fbitsP)) f |= WORK_P4;
      if (a.t3.present && filter_pass2(f3, (lo32 & 0xFFFFFFu) * HASH_A, a.t3.fbits)) f |= WORK_3;
      if (TINY) {
        if (a.has2 && a.tiny2[lo32 & 0xFFFFu]) f |= WORK_2;
        if (a.has1 && a.tiny1[lo32 & 0xFFu]) f |= WORK_1;
      }
      if (rel < 0 || rel >= (int)span) f = 0;
      const unsigned m = __ballot_sync(FULL, f != 0);
      if (f) work[n_work + __popc(m & lt)] = ((uint32_t)rel << 4) | f;
      n_work += __popc(m);
    }
    if (n_work > SM_WORK - 128) {
      __syncwarp();
      drain_work(a, lane, B0, work, n_work, s_start, s_nc, cand);
      n_work = 0;
    }
  }
  __syncwarp();
  drain_work(a, lane, B0, work, n_work, s_start, s_nc, cand);
}

// phase 2, whole warp: `nc` = this lane's candidate count (<= SM_CANDBUF), s / len = this lane's name.
// Matches go to hits[owner][...] with n_hits[owner] counting them (may run past SM_HITBUF: overflow).
__device__ __forceinline__ void resolve_candidates(const SearchArgs& a, int lane, uint32_t nc, unsigned long long s,
                                                   uint32_t len, const uint16_t (*cand)[SM_CANDBUF],
                                                   uint16_t (*hits)[SM_HITBUF], uint32_t* n_hits, uint16_t* pref) {
  const uint32_t inc = warp_incl_scan_u32(nc);
  const uint32_t total = __shfl_sync(FULL, inc, 31);
  pref[lane] = (uint16_t)(inc - nc);
  __syncwarp();
  for (uint32_t j0 = 0; j0 < total; j0 += 32) {
    const uint32_t j = j0 + lane;
    const bool valid = j < total;
    uint32_t owner = 0;   // last lane whose first candidate is <= j
    if (valid) {
#pragma unroll
      for (int st = 16; st > 0; st >>= 1)
        if (pref[owner + st] <= j) owner += st;
    }
    const unsigned long long so = __shfl_sync(FULL, s, owner);
    const uint32_t leno = __shfl_sync(FULL, len, owner);
    if (!valid) continue;
    const uint32_t c = cand[owner][j - pref[owner]];
    const uint32_t cls = c >> CAND_POS_BITS, p = c & ((1u << CAND_POS_BITS) - 1u);
    const uint32_t r = leno - p;
    // 16 window bytes at so + p
    const unsigned long long at = so + p;
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(a.lower + (at & ~3ull));
    const uint32_t sh = (uint32_t)(at & 3ull) * 8;
    const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3), w4 = __ldg(wp + 4);
    const uint32_t lo32 = sh ? __funnelshift_r(w0, w1, sh) : w0, hi32 = sh ? __funnelshift_r(w1, w2, sh) : w1;
    const uint32_t lo2 = sh ? __funnelshift_r(w2, w3, sh) : w2, hi2 = sh ? __funnelshift_r(w3, w4, sh) : w3;
    const unsigned long long k8 = ((unsigned long long)hi32 << 32) | lo32;
    auto hit = [&](uint32_t u) {
      const uint32_t slot = atomicAdd(&n_hits[owner], 1u);
      if (slot < SM_HITBUF) hits[owner][slot] = (uint16_t)u;
    };
    if (cls >= 5) {
      const uint32_t u1 = cls == 5 ? a.tiny2[lo32 & 0xFFFFu] : a.tiny1[lo32 & 0xFFu];
      if (u1) hit(u1 - 1);
      continue;
    }
    const SlotTable& tb = cls == 0 ? a.t16 : cls == 1 ? a.t12 : cls == 2 ? a.t8 : cls == 3 ? a.t4 : a.t3;
    const KeyFp kf = class_key((int)cls, lo32, cls >= 3 ? 0u : hi32, cls >= 2 ? 0u : lo2, cls >= 1 ? 0u : hi2);
    for (uint32_t u = table_find(tb, kf.key, kf.fp); u != PAT_NIL;) {
      const PatInfo pi = load_pat(a.pat, u);
      if (pi.len <= r && (k8 & low_bytes_mask(pi.len)) == pi.pk8) {   // first min(len, 8) bytes
        bool ok = true;
        if (pi.len > 8) {
          // bytes 8.. four at a time: pattern words are 4-byte aligned (poff is), name words come from the
          // aligned pair around them; bytes 8..15 are still in the window registers
          const uint32_t* pw = reinterpret_cast<const uint32_t*>(a.pbytes + a.poff[u]);
          for (uint32_t t = 8; t < pi.len && ok; t += 4) {
            const uint32_t k = t >> 2;
            const uint32_t nw = k == 2 ? lo2 : (k == 3 ? hi2 : (sh ? __funnelshift_r(__ldg(wp + k), __ldg(wp + k + 1), sh) : __ldg(wp + k)));
            const uint32_t rem = pi.len - t;
            const uint32_t mask = rem >= 4 ? 0xFFFFFFFFu : ((1u << (8 * rem)) - 1u);
            ok = ((nw ^ __ldg(pw + k)) & mask) == 0;
          }
        }
        if (ok) hit(u);
      }
      u = pi.next;
    }
  }
  __syncwarp();
}

// every lane on its own list: keeps one copy of each pattern; returns the list length
__device__ __forceinline__ uint32_t dedup_hits(uint16_t* mine, uint32_t n) {
  for (uint32_t t = 1; t < n;) {
    const uint16_t u = mine[t];
    bool dup = false;
    for (uint32_t t2 = 0; t2 < t && !dup; ++t2) dup = mine[t2] == u;
    if (dup) mine[t] = mine[--n];
    else ++t;
  }
  return n;
}

// every pattern against one name, by the whole warp (names whose list overflowed)
template <class F>
__device__ __forceinline__ void exhaustive_name(const SearchArgs& a, uint32_t id, int lane, F deliver) {
  const unsigned long long s = a.name_off[id], e = a.name_off[id + 1];
  const uint32_t nn = (uint32_t)(e - s);
  for (uint32_t u = lane; u < a.U; u += 32) {
    const uint32_t len = a.pat[u].len;
    bool hit = len == 0;
    const uint8_t* pb = a.pbytes + a.poff[u];
    for (uint32_t p = 0; !hit && p + len <= nn; ++p) {
      uint32_t t = 0;
      while (t < len && a.lower[s + p + t] == pb[t]) ++t;
      hit = t == len;
    }
    if (hit) deliver(u, id);
  }
}

// One lane per name: the 32 lanes of a warp scan 32 consecutive names in lock step, so the
// (frequent) hit path runs with many lanes active.  A lane collects the distinct patterns of its
// name in shared memory; the warp then handles the 32 lists in lane order == name order, which keeps
// every row ascending.
//   MODE 0  count: per (pattern, tile) counters, and the lists are appended to a hit log so that
//           the emit pass does not have to match again
//   MODE 1  emit: streams the log (32 hits per warp step; hits of one pattern inside a step get
//           consecutive slots through __match_any_sync), re-matching only the steps that were
//           not logged (log full, or a name with more than SM_HITBUF patterns)
//   MODE 2  first: atomicMin per pattern
template <int MODE>
__global__ void __launch_bounds__(SM_THREADS) search_multi_kernel(const SearchArgs a) {
  // dynamic shared memory: the two filters every position asks (4-byte prefix filter, 3-byte filter), then
  // the warps' hit lists and candidate queues
  extern __shared__ __align__(16) uint32_t sm_u32[];
  uint32_t* fP = sm_u32;
  uint32_t* f3 = fP + a.filt_wordsP;
  __shared__ uint32_t s_nh[SM_WARPS][32];
  __shared__ uint16_t s_pref[SM_WARPS][32];   // (u16 / the sizes above: four CTAs of config 3 fit an SM by 128 bytes)
  __shared__ uint32_t s_start[SM_WARPS][36];
  __shared__ uint32_t s_nc[SM_WARPS][32];
  __shared__ uint32_t s_work[SM_WARPS][SM_WORK];
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  const uint32_t fw = a.filt_wordsP + a.filt_words3;
  typedef uint16_t HitLists[32][SM_HITBUF];
  typedef uint16_t CandQueues[32][SM_CANDBUF];
  HitLists* const s_hits = reinterpret_cast<HitLists*>(sm_u32 + ((fw + 3u) & ~3u));
  CandQueues* const s_cand = reinterpret_cast<CandQueues*>(s_hits + SM_WARPS);
  {
    const uint32_t* gP = a.filt + a.filt_words16 + a.filt_words12 + a.filt_words8 + a.filt_words4 + a.filt_words3;
    const uint32_t* g3 = a.filt + a.filt_words16 + a.filt_words12 + a.filt_words8 + a.filt_words4;
    for (uint32_t i = tid; i < a.filt_wordsP; i += SM_THREADS) fP[i] = gP[i];
    for (uint32_t i = tid; i < a.filt_words3; i += SM_THREADS) f3[i] = g3[i];
  }
  __syncthreads();

  const uint32_t tile = blockIdx.x * SM_WARPS + wib;
  if (tile >= a.n_tiles) return;
  const uint32_t n0 = a.v_lo + tile * a.per_tile;
  const uint32_t n1 = min(a.v_hi, n0 + a.per_tile);
  if (MODE == 1 && !a.tile_flag[tile]) return;   // every step of the tile is in the log: search_emit_log_kernel
  uint32_t* my_cnt = (MODE == 2) ? nullptr : a.cnt + tile;   // element of pattern u: my_cnt[u * n_tiles]
  uint16_t* mine = s_hits[wib][lane];

  // a hit goes to the row of the FIRST query that asked for the pattern; rows of queries repeating a
  // text are copied from it afterwards (search_replicate_kernel) -- config 3 asks its short, high-hit
  // substrings up to 15 times each: two thirds of the 250 M output entries are such copies
  auto write_rows = [&](uint32_t u, uint32_t id, uint32_t local) {
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(a.prow + u));
    const unsigned long long pos = (((unsigned long long)raw.y << 32) | raw.x) + local;
    if (pos < a.cap) a.out_ids[pos] = id;
  };
  auto deliver = [&](uint32_t u, uint32_t id) {
    if (MODE == 0) atomicAdd(&my_cnt[(size_t)u * a.n_tiles], 1u);
    else if (MODE == 1) write_rows(u, id, atomicAdd(&my_cnt[(size_t)u * a.n_tiles], 1u));
    else atomicMin(&a.first[u], id);
  };

  for (uint32_t i0 = n0; i0 < n1; i0 += 32) {
    const size_t step = (size_t)tile * a.steps_per_tile + (i0 - n0) / 32;
    if (MODE == 1) {
      const uint32_t logged = a.step_cnt[step];
      if (logged != 0xFFFFFFFFu) {
        const unsigned long long b = a.step_base[step];
        for (uint32_t j0 = 0; j0 < logged; j0 += 32) {
          const uint32_t j = j0 + lane;
          const bool valid = j < logged;
          const uint32_t ent = valid ? a.log[b + j] : 0u;
          const uint32_t u = ent & ((1u << LOG_U_BITS) - 1u);
          const unsigned peers = __match_any_sync(FULL, valid ? u : (0x10000u + lane));
          const int leader = __ffs(peers) - 1;
          uint32_t base = 0;
          if (valid && lane == leader) base = atomicAdd(&my_cnt[(size_t)u * a.n_tiles], (uint32_t)__popc(peers));
          base = __shfl_sync(FULL, base, leader);
          if (valid) write_rows(u, i0 + ((ent >> LOG_U_BITS) & 31u), base + __popc(peers & lanemask_lt()));
        }
        continue;
      }
    }
    uint32_t nh;
    {
      const uint32_t i = i0 + lane;
      unsigned long long s = 0;
      uint32_t len = 0;
      if (i < n1) {
        s = a.name_off[i];
        len = (uint32_t)min(a.name_off[i + 1] - s, 0xFFFFFFFFull);
      }
      // the step's byte range and the names' starts inside it
      const uint32_t n_valid = min(32u, n1 - i0);
      const unsigned long long B0 = __shfl_sync(FULL, s, 0);
      const unsigned long long B1 = __shfl_sync(FULL, s + len, (int)n_valid - 1);
      const bool spans = B1 - B0 < (1ull << 27);   // a parked position is 28 bits
      s_start[wib][lane] = (uint32_t)((i < n1 ? s : B1) - B0);
      if (lane == 0) s_start[wib][32] = (uint32_t)(B1 - B0);
      // names too long for a 13-bit position (or a step too long for 32-bit offsets) take the exhaustive path
      s_nc[wib][lane] = (len >= (1u << CAND_POS_BITS) || !spans) ? (uint32_t)SM_CANDBUF + 1u : 0u;
      __syncwarp();
      if (spans) {
        if (a.has2 | a.has1) scan_step<true>(a, fP, f3, lane, B0, B1, s_start[wib], s_nc[wib], s_cand[wib], s_work[wib]);
        else scan_step<false>(a, fP, f3, lane, B0, B1, s_start[wib], s_nc[wib], s_cand[wib], s_work[wib]);
      }
      const uint32_t nc = i < n1 ? s_nc[wib][lane] : 0u;
      const bool cand_overflow = nc > SM_CANDBUF;
      // the empty pattern is contained in every name (CCS:1513: "".contains)
      if (i < n1 && a.empty_plus1) mine[0] = (uint16_t)(a.empty_plus1 - 1);
      s_nh[wib][lane] = (i < n1 && a.empty_plus1) ? 1u : 0u;
      __syncwarp();
      resolve_candidates(a, lane, cand_overflow ? 0u : nc, s, len, s_cand[wib], s_hits[wib], s_nh[wib], s_pref[wib]);
      const uint32_t raw_n = s_nh[wib][lane];
      nh = (cand_overflow || raw_n > SM_HITBUF) ? SM_HITBUF + 1 : dedup_hits(mine, raw_n);
    }
    __syncwarp();
    s_nh[wib][lane] = nh;
    __syncwarp();
    if (MODE == 0) {
      // Counters and hit log in one sweep over the 32 lists laid back to back (lane order == name order),
      // 32 hits per step with every lane busy: hits of one pattern inside a step take consecutive ranks
      // from ONE add on the (pattern, tile) counter (__match_any_sync), and the rank goes into the log
      // entry, so the emit pass places a logged hit without any cursor of its own.
      const bool overflow = __any_sync(FULL, nh > SM_HITBUF);
      if (!overflow) {
        const uint32_t inc = warp_incl_scan_u32(nh);
        const uint32_t tot = __shfl_sync(FULL, inc, 31);
        s_pref[wib][lane] = (uint16_t)(inc - nh);
        unsigned long long b = 0;
        if (lane == 0 && tot) b = atomicAdd(a.log_used, (unsigned long long)tot);
        b = __shfl_sync(FULL, b, 0);
        bool fits = b + tot <= a.log_cap;
        __syncwarp();
        for (uint32_t e0 = 0; e0 < tot; e0 += 32) {
          const uint32_t e = e0 + lane;
          const bool valid = e < tot;
          uint32_t owner = 0, u = 0;
          if (valid) {
#pragma unroll
            for (int st = 16; st > 0; st >>= 1)
              if (s_pref[wib][owner + st] <= e) owner += st;
            u = s_hits[wib][owner][e - s_pref[wib][owner]];
          }
          const unsigned peers = __match_any_sync(FULL, valid ? u : (0x10000u + lane));
          const int leader = __ffs(peers) - 1;
          uint32_t base = 0;
          if (valid && lane == leader) base = atomicAdd(&my_cnt[(size_t)u * a.n_tiles], (uint32_t)__popc(peers));
          base = __shfl_sync(FULL, base, leader);
          const uint32_t idx = base + __popc(peers & lanemask_lt());
          if (__any_sync(FULL, valid && idx >= LOG_MAX_TILE)) fits = false;   // cannot happen while per_tile <= LOG_MAX_TILE
          if (fits && valid) a.log[b + e] = log_pack(idx, owner, u);
        }
        if (lane == 0) {
          a.step_base[step] = b;
          a.step_cnt[step] = fits ? tot : 0xFFFFFFFFu;
          if (!fits) a.tile_flag[tile] = 1;
        }
        __syncwarp();
        continue;
      }
      if (lane == 0) {   // a name with more patterns than a list holds: the step is matched again by the emit pass
        a.step_base[step] = 0;
        a.step_cnt[step] = 0xFFFFFFFFu;
        a.tile_flag[tile] = 1;
      }
    }
    // counters / rows / minima, in lane order == name order
    for (int l = 0; l < 32; ++l) {
      const uint32_t n_l = s_nh[wib][l];
      const uint32_t id = i0 + l;
      if (n_l == 0) continue;
      if (n_l <= SM_HITBUF) {
        for (uint32_t j = lane; j < n_l; j += 32) deliver(s_hits[wib][l][j], id);
      } else {
        exhaustive_name(a, id, lane, deliver);
      }
    }
    __syncwarp();
  }
}

// Emit for tiles whose every step is in the log: a pure stream.  One warp per step; an entry knows its
// pattern, its name and its rank among the tile's hits of the pattern, so
//     position = row of the pattern's first query + hits of the pattern in earlier tiles + rank
// needs no cursor, no atomics and no order between entries.
__global__ void __launch_bounds__(256) search_emit_log_kernel(const SearchArgs a) {
  const int lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
  const size_t n_steps = (size_t)a.n_tiles * a.steps_per_tile;
  for (size_t step = warp; step < n_steps; step += n_warps) {
    const uint32_t tile = (uint32_t)(step / a.steps_per_tile);
    if (a.tile_flag[tile]) continue;
    const uint32_t in_tile = (uint32_t)(step - (size_t)tile * a.steps_per_tile);
    if (in_tile * 32u >= a.per_tile) continue;
    const uint32_t i0 = a.v_lo + tile * a.per_tile + in_tile * 32u;
    if (i0 >= a.v_hi) continue;
    const uint32_t logged = a.step_cnt[step];
    const unsigned long long b = a.step_base[step];
    const uint32_t* __restrict__ tcnt = a.cnt + tile;
    for (uint32_t j = lane; j < logged; j += 32) {
      const uint32_t ent = __ldg(&a.log[b + j]);
      const uint32_t u = ent & ((1u << LOG_U_BITS) - 1u);
      const uint2 raw = __ldg(reinterpret_cast<const uint2*>(a.prow + u));
      const unsigned long long pos = (((unsigned long long)raw.y << 32) | raw.x) + tcnt[(size_t)u * a.n_tiles] + (ent >> LOG_IDX_SHIFT);
      if (pos < a.cap) a.out_ids[pos] = i0 + ((ent >> LOG_U_BITS) & 31u);
    }
  }
}

// exclusive scan of every pattern's tile counts (one warp per pattern, the pattern's tiles are
// contiguous); tot[u] = row length
__global__ void __launch_bounds__(256) search_colscan_kernel(uint32_t* cnt, uint32_t n_tiles, uint32_t U,
                                                             uint32_t* tot) {
  const int lane = threadIdx.x & 31;
  const uint32_t u = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (u >= U) return;
  uint32_t* row = cnt + (size_t)u * n_tiles;
  uint32_t run = 0;
  for (uint32_t t0 = 0; t0 < n_tiles; t0 += 32) {
    const uint32_t t = t0 + lane;
    const uint32_t c = t < n_tiles ? row[t] : 0;
    const uint32_t inc = warp_incl_scan_u32(c);
    if (t < n_tiles) row[t] = run + inc - c;
    run += __shfl_sync(FULL, inc, 31);
  }
  if (lane == 0) tot[u] = run;
}

// once out_off is known: where each pattern's hits go
__global__ void search_rows_kernel(const uint32_t* __restrict__ uq_off, const uint32_t* __restrict__ uq,
                                   const unsigned long long* __restrict__ out_off, uint32_t U, PatRow* prow) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  PatRow r;
  r.uq_begin = uq_off[u];
  r.n_queries = uq_off[u + 1] - r.uq_begin;
  r.row_base = r.n_queries ? out_off[uq[r.uq_begin]] : 0;
  prow[u] = r;
}

// Row of a query whose text an earlier query of the batch already asked: a copy of that query's row.
// grid = (duplicate queries, slices); a block copies every gridDim.y-th 4096-entry piece of its row.
__global__ void __launch_bounds__(256) search_replicate_kernel(const uint32_t* __restrict__ dup_q,
                                                               const uint32_t* __restrict__ dup_first,
                                                               const unsigned long long* __restrict__ out_off,
                                                               uint32_t* out_ids, unsigned long long cap) {
  const uint32_t q = dup_q[blockIdx.x], f = dup_first[blockIdx.x];
  const unsigned long long src = out_off[f], dst = out_off[q], n = out_off[f + 1] - src;
  for (unsigned long long p0 = (unsigned long long)blockIdx.y * 4096; p0 < n; p0 += (unsigned long long)gridDim.y * 4096) {
    const unsigned long long p1 = min(n, p0 + 4096);
    for (unsigned long long p = p0 + threadIdx.x; p < p1; p += blockDim.x)
      if (dst + p < cap && src + p < cap) out_ids[dst + p] = out_ids[src + p];
  }
}

__global__ void search_query_counts_kernel(const uint32_t* __restrict__ tot, const uint32_t* __restrict__ q2u,
                                           unsigned long long nq, uint32_t* qcnt) {
  const unsigned long long q = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (q < nq) {
    const uint32_t u = q2u[q];
    qcnt[q] = (u == 0xFFFFFFFFu) ? 0u : tot[u];
  }
}

__global__ void search_first_rows_kernel(const uint32_t* __restrict__ first, const uint32_t* __restrict__ q2u,
                                         unsigned long long nq, uint32_t* qcnt) {
  const unsigned long long q = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (q < nq) qcnt[q] = first[q2u[q]] != 0xFFFFFFFFu;
}
__global__ void search_first_emit_kernel(const uint32_t* __restrict__ first, const uint32_t* __restrict__ q2u,
                                         unsigned long long nq, const unsigned long long* __restrict__ out_off,
                                         uint32_t* out_ids, unsigned long long cap) {
  const unsigned long long q = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (q < nq) {
    const uint32_t f = first[q2u[q]];
    if (f != 0xFFFFFFFFu && out_off[q] < cap) out_ids[out_off[q]] = f;
  }
}

// ---------------------------------------------------------------------------------------------
// host side: pattern tables
// ---------------------------------------------------------------------------------------------
struct HostTables {
  std::vector<std::string> pats;                 // distinct patterns
  std::vector<uint32_t> q2u;
  std::vector<PatSlot> slot[5];                  // classes 16, 12, 8, 4, 3
  std::vector<PatInfo> pat;
  std::vector<uint32_t> filt[5];
  uint32_t fbits[5] = {0, 0, 0, 0, 0};
  std::vector<uint32_t> filtP;                   // 4-byte prefixes of every pattern of >= 4 bytes
  uint32_t fbitsP = 12, presentP = 0;
  std::vector<uint32_t> poff;
  std::vector<uint8_t> pbytes;
  std::vector<uint32_t> tiny2, tiny1;
  uint32_t empty_plus1 = 0;
  std::vector<uint32_t> uq_off, uq;
};

static void build_tables(uint64_t nq, const uint32_t* q_off, const uint8_t* q_bytes, HostTables& T) {
  std::unordered_map<std::string, uint32_t> ids;
  T.q2u.resize(nq);
  for (uint64_t q = 0; q < nq; ++q) {
    std::string p(reinterpret_cast<const char*>(q_bytes) + q_off[q], q_off[q + 1] - q_off[q]);
    auto it = ids.find(p);
    if (it == ids.end()) {
      it = ids.emplace(p, (uint32_t)T.pats.size()).first;
      T.pats.push_back(p);
    }
    T.q2u[q] = it->second;
  }
  const uint32_t U = (uint32_t)T.pats.size();
  {
    // short patterns hit most often: give them the low ids so that the cells of the cursor
    // matrix that are touched all the time sit together (and stay in L2)
    std::vector<uint32_t> order(U), rank(U);
    for (uint32_t u = 0; u < U; ++u) order[u] = u;
    std::stable_sort(order.begin(), order.end(),
                     [&](uint32_t x, uint32_t y) { return T.pats[x].size() < T.pats[y].size(); });
    std::vector<std::string> sorted(U);
    for (uint32_t r = 0; r < U; ++r) {
      rank[order[r]] = r;
      sorted[r] = std::move(T.pats[order[r]]);
    }
    T.pats.swap(sorted);
    for (uint64_t q = 0; q < nq; ++q) T.q2u[q] = rank[T.q2u[q]];
  }
  T.pat.resize(U);
  T.poff.resize(U + 1);
  T.tiny2.assign(65536, 0);
  T.tiny1.assign(256, 0);
  struct KeyedPat {
    unsigned long long key;
    uint32_t fp, u, raw;
    bool operator<(const KeyedPat& o) const { return key != o.key ? key < o.key : (fp != o.fp ? fp < o.fp : u < o.u); }
  };
  std::vector<KeyedPat> cls[5];
  for (uint32_t u = 0; u < U; ++u) {
    const std::string& p = T.pats[u];
    T.pat[u].len = (uint32_t)p.size();
    T.pat[u].next = PAT_NIL;
    T.poff[u] = (uint32_t)T.pbytes.size();   // a multiple of 4: the verification compares words
    T.pbytes.insert(T.pbytes.end(), p.begin(), p.end());
    T.pbytes.resize((T.pbytes.size() + 3) & ~size_t(3), 0);
    unsigned long long k = 0;
    for (size_t i = 0; i < p.size() && i < 8; ++i) k |= (unsigned long long)(uint8_t)p[i] << (8 * i);
    T.pat[u].pk8 = k;
    if (p.empty()) T.empty_plus1 = u + 1;
    else if (p.size() == 1) T.tiny1[k & 0xFF] = u + 1;
    else if (p.size() == 2) T.tiny2[k & 0xFFFF] = u + 1;
    else {
      unsigned long long k2 = 0;
      for (size_t i = 8; i < p.size() && i < 16; ++i) k2 |= (unsigned long long)(uint8_t)p[i] << (8 * (i - 8));
      const int c = p.size() >= 16 ? 0 : p.size() >= 12 ? 1 : p.size() >= 8 ? 2 : p.size() >= 4 ? 3 : 4;
      const KeyFp kf = class_key(c, (uint32_t)k, (uint32_t)(k >> 32), (uint32_t)k2, c == 0 ? (uint32_t)(k2 >> 32) : 0u);
      cls[c].push_back({kf.key, kf.fp, u, kf.raw});
    }
  }
  T.poff[U] = (uint32_t)T.pbytes.size();
  T.pbytes.resize(T.pbytes.size() + 16, 0);
  for (int c = 0; c < 5; ++c) {
    auto& v = cls[c];
    std::sort(v.begin(), v.end());
    size_t nkeys = 0;
    for (size_t i = 0; i < v.size(); ++i) nkeys += (i == 0 || v[i].key != v[i - 1].key || v[i].fp != v[i - 1].fp);
    uint32_t size = 16;
    while (size < 2 * nkeys + 1) size <<= 1;
    PatSlot empty;
    empty.key = 0;
    empty.first = PAT_NIL;
    empty.fp = 0;
    T.slot[c].assign(size, empty);
    uint32_t fb = 12;
    while ((1ull << fb) < 32 * nkeys && fb < (c == 0 ? 17u : 15u)) ++fb;
    T.fbits[c] = fb;
    T.filt[c].assign((1u << fb) / 32, 0);
    for (size_t i = 0; i < v.size();) {
      size_t j = i;
      while (j < v.size() && v[j].key == v[i].key && v[j].fp == v[i].fp) ++j;
      uint32_t h = hash_slot(v[i].fp, size - 1);
      while (T.slot[c][h].first != PAT_NIL) h = (h + 1) & (size - 1);
      T.slot[c][h].key = v[i].key;
      T.slot[c][h].first = v[i].u;
      T.slot[c][h].fp = v[i].fp;
      for (size_t t = i; t + 1 < j; ++t) T.pat[v[t].u].next = v[t + 1].u;   // patterns sharing the key
      const uint32_t fi = hash_filter(v[i].raw, fb), fj = hash_filter2(v[i].raw, fb);
      T.filt[c][fi >> 5] |= 1u << (fi & 31);
      T.filt[c][fj >> 5] |= 1u << (fj & 31);
      i = j;
    }
  }
  {
    // the filter every text position asks first: is there ANY pattern of >= 4 bytes starting with these 4?
    std::vector<uint32_t> pre;
    for (int c = 0; c < 4; ++c)
      for (const KeyedPat& kp : cls[c]) pre.push_back((uint32_t)T.pat[kp.u].pk8);
    std::sort(pre.begin(), pre.end());
    pre.erase(std::unique(pre.begin(), pre.end()), pre.end());
    uint32_t fb = 12;
    while ((1ull << fb) < 32 * pre.size() && fb < 17u) ++fb;
    T.fbitsP = fb;
    T.presentP = !pre.empty();
    T.filtP.assign((1u << fb) / 32, 0);
    for (uint32_t k4 : pre) {
      const uint32_t raw = class_raw(3, k4, 0, 0, 0);
      const uint32_t fi = hash_filter(raw, fb), fj = hash_filter2(raw, fb);
      T.filtP[fi >> 5] |= 1u << (fi & 31);
      T.filtP[fj >> 5] |= 1u << (fj & 31);
    }
  }
  // pattern -> queries
  T.uq_off.assign(U + 1, 0);
  for (uint64_t q = 0; q < nq; ++q) ++T.uq_off[T.q2u[q] + 1];
  for (uint32_t u = 0; u < U; ++u) T.uq_off[u + 1] += T.uq_off[u];
  T.uq.resize(nq);
  std::vector<uint32_t> cur(T.uq_off.begin(), T.uq_off.end() - 1);
  for (uint64_t q = 0; q < nq; ++q) T.uq[cur[T.q2u[q]]++] = (uint32_t)q;
}

template <class T>
static int32_t put(CallCtx* c, DevBuf& arena, size_t* used, const std::vector<T>& v, const T** out) {
  const size_t bytes = std::max<size_t>(v.size() * sizeof(T), 16);
  const size_t at = (*used + 255) & ~size_t(255);
  *out = reinterpret_cast<const T*>(static_cast<char*>(arena.p) + at);
  if (!v.empty())
    DMG_CUDA(cudaMemcpyAsync(const_cast<T*>(*out), v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  *used = at + bytes;
  return DMG_OK;
}

// Runs the multi-pattern search for kind CONTAINS_ALL / CONTAINS_FIRST on device buffers of the
// context; leaves out_off / out_ids on the device and the total in c->h_counters->total.
int32_t run_search_multi(CallCtx* c, dmg_graph* g, int kind, uint32_t v_lo, uint32_t v_hi, uint64_t nq,
                         const uint32_t* q_off, const uint8_t* q_bytes, uint64_t dev_cap, bool sizes_only) {
  Engine* e = c->eng;
  HostTables T;
  build_tables(nq, q_off, q_bytes, T);
  const uint32_t U = (uint32_t)T.pats.size();
  if (U > SM_MAX_PATTERNS && kind == DMG_SEARCH_CONTAINS_ALL) {
    set_error("dmg_search_batch: %u distinct queries in one batch; split it into batches of <= %u", U, SM_MAX_PATTERNS);
    return DMG_EINVAL;
  }
  // one arena for every table
  size_t need = 4096;
  auto sz = [&](size_t n, size_t s) { need += ((n * s + 255) & ~size_t(255)) + 256; };
  for (int k = 0; k < 5; ++k) {
    sz(T.slot[k].size(), 16);
    sz(T.filt[k].size(), 4);
  }
  sz(T.filtP.size(), 4);
  sz(U, 16); sz(U, 16); sz(3 * 65536, 4); sz(U + 1, 4); sz(T.pbytes.size(), 1);
  sz(65536, 4); sz(256, 4); sz(U + 1, 4); sz(nq, 4); sz(nq, 4); sz(nq, 4); sz(nq, 4);
  DMG_TRY(c->tmp0.reserve(need));
  size_t used = 0;
  SearchArgs a;
  memset(&a, 0, sizeof(a));
  SlotTable* st[5] = {&a.t16, &a.t12, &a.t8, &a.t4, &a.t3};
  std::vector<uint32_t> filt_all;
  for (int k = 0; k < 5; ++k) {
    DMG_TRY(put(c, c->tmp0, &used, T.slot[k], &st[k]->slot));
    st[k]->mask = (uint32_t)T.slot[k].size() - 1;
    st[k]->fbits = T.fbits[k];
    st[k]->present = 0;
    for (const PatSlot& x : T.slot[k]) st[k]->present |= x.first != PAT_NIL;
    filt_all.insert(filt_all.end(), T.filt[k].begin(), T.filt[k].end());
  }
  a.filt_words16 = (uint32_t)T.filt[0].size();
  a.filt_words12 = (uint32_t)T.filt[1].size();
  a.filt_words8 = (uint32_t)T.filt[2].size();
  a.filt_words4 = (uint32_t)T.filt[3].size();
  a.filt_words3 = (uint32_t)T.filt[4].size();
  a.filt_wordsP = (uint32_t)T.filtP.size();
  a.fbitsP = T.fbitsP;
  a.presentP = T.presentP;
  filt_all.insert(filt_all.end(), T.filtP.begin(), T.filtP.end());
  DMG_TRY(put(c, c->tmp0, &used, filt_all, &a.filt));
  DMG_TRY(put(c, c->tmp0, &used, T.pat, &a.pat));
  {
    std::vector<PatRow> rows(U);
    const PatRow* prow_c = nullptr;
    DMG_TRY(put(c, c->tmp0, &used, rows, &prow_c));   // space only; search_rows_kernel fills it
    a.prow = const_cast<PatRow*>(prow_c);
  }
  DMG_TRY(put(c, c->tmp0, &used, T.poff, &a.poff));
  DMG_TRY(put(c, c->tmp0, &used, T.pbytes, &a.pbytes));
  DMG_TRY(put(c, c->tmp0, &used, T.tiny2, &a.tiny2));
  DMG_TRY(put(c, c->tmp0, &used, T.tiny1, &a.tiny1));
  DMG_TRY(put(c, c->tmp0, &used, T.uq_off, &a.uq_off));
  DMG_TRY(put(c, c->tmp0, &used, T.uq, &a.uq));
  const uint32_t* d_q2u = nullptr;
  DMG_TRY(put(c, c->tmp0, &used, T.q2u, &d_q2u));
  // queries repeating the text of an earlier query (their rows are copies)
  std::vector<uint32_t> dup_q, dup_first;
  for (uint32_t u = 0; u < U; ++u)
    for (uint32_t t = T.uq_off[u] + 1; t < T.uq_off[u + 1]; ++t) {
      dup_q.push_back(T.uq[t]);
      dup_first.push_back(T.uq[T.uq_off[u]]);
    }
  const uint32_t n_dup = (uint32_t)dup_q.size();
  const uint32_t *d_dup_q = nullptr, *d_dup_first = nullptr;
  DMG_TRY(put(c, c->tmp0, &used, dup_q, &d_dup_q));
  DMG_TRY(put(c, c->tmp0, &used, dup_first, &d_dup_first));
  for (uint32_t x : T.tiny2) a.has2 |= x != 0;
  for (uint32_t x : T.tiny1) a.has1 |= x != 0;
  a.empty_plus1 = T.empty_plus1;
  a.U = U;
  a.name_off = reinterpret_cast<const unsigned long long*>(g->d_name_off);
  a.lower = g->d_lower;
  a.lower_size = g->info.name_bytes + 32;
  a.v_lo = v_lo;
  a.v_hi = v_hi;
  c->stats.h2d_bytes += used;

  const uint32_t range = v_hi - v_lo;
  const size_t smem = (((size_t)a.filt_wordsP + a.filt_words3 + 3) & ~size_t(3)) * 4 +
                      (size_t)SM_WARPS * 32 * (SM_HITBUF + SM_CANDBUF) * sizeof(uint16_t);
  {
    DMG_CUDA(cudaFuncSetAttribute(search_multi_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DMG_CUDA(cudaFuncSetAttribute(search_multi_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DMG_CUDA(cudaFuncSetAttribute(search_multi_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  int occ = 1;
  DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, search_multi_kernel<0>, SM_THREADS, smem));
  // tiles: one per resident warp, but not smaller than 64 names and bounded by the matrix budget
  uint64_t tiles = (uint64_t)e->sm_count * std::max(occ, 1) * SM_WARPS;
  tiles = std::min<uint64_t>(tiles, ((uint64_t)range + 63) / 64);
  if (kind == DMG_SEARCH_CONTAINS_ALL)
    tiles = std::min<uint64_t>(tiles, std::max<uint64_t>(1, (1ull << 30) / (4ull * std::max<uint32_t>(U, 1))));
  tiles = std::max<uint64_t>(tiles, 1);
  // a log entry has 13 bits for a hit's rank inside its tile (the matrix budget above wins on a huge range:
  // such tiles then simply fall back to the matching emit pass)
  if (kind == DMG_SEARCH_CONTAINS_ALL)
    tiles = std::max<uint64_t>(tiles, std::min<uint64_t>(((uint64_t)range + LOG_MAX_TILE - 1) / LOG_MAX_TILE,
                                                         std::max<uint64_t>(1, (1ull << 30) / (4ull * std::max<uint32_t>(U, 1)))));
  a.n_tiles = (uint32_t)tiles;
  a.per_tile = (uint32_t)((range + tiles - 1) / tiles);
  a.steps_per_tile = (a.per_tile + 31) / 32;
  {
    // hit log of the count pass: the emit pass streams it instead of matching again.  Sized for 32
    // hits per name; steps that do not fit are simply matched again by the emit pass.
    const size_t n_steps = (size_t)tiles * std::max<uint32_t>(a.steps_per_tile, 1);
    DMG_TRY(c->tmp3.reserve(n_steps * 12 + 64 + (size_t)tiles * 4));
    a.step_base = c->tmp3.as<unsigned long long>();
    a.step_cnt = reinterpret_cast<uint32_t*>(a.step_base + n_steps);
    a.tile_flag = a.step_cnt + ((n_steps + 3) & ~size_t(3));
    DMG_CUDA(cudaMemsetAsync(a.tile_flag, 0, (size_t)tiles * 4, c->stream));
    a.log_used = &c->d_counters->aux[3];
    a.log_cap = 0;
    if (kind == DMG_SEARCH_CONTAINS_ALL && !sizes_only) {
      const uint64_t want = std::min<uint64_t>(e->scratch_budget / 4, std::max<uint64_t>(1u << 20, 32ull * range));
      DMG_TRY(c->scratch.reserve(want * 4));
      a.log = c->scratch.as<uint32_t>();
      a.log_cap = want;
    }
  }
  const unsigned blocks = (unsigned)((tiles + SM_WARPS - 1) / SM_WARPS);

  DMG_TRY(c->out_off.reserve((nq + 1) * 8));
  DMG_TRY(c->tmp2.reserve(nq * 4 + (size_t)U * 4 + 512));   // per-query counts | per-pattern totals/first
  uint32_t* qcnt = c->tmp2.as<uint32_t>();
  uint32_t* tot = qcnt + ((nq + 63) & ~63ull);
  const uint64_t scan_tiles = (nq + 2047) / 2048;
  DMG_TRY(c->status.reserve(scan_tiles * 8));
  DMG_CUDA(cudaMemsetAsync(c->status.p, 0, scan_tiles * 8, c->stream));
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  a.out_off = c->out_off.as<unsigned long long>();
  a.cap = dev_cap;
  const unsigned qb = (unsigned)((nq + 255) / 256);

  if (kind == DMG_SEARCH_CONTAINS_FIRST) {
    a.first = tot;
    DMG_CUDA(cudaMemsetAsync(tot, 0xFF, (size_t)U * 4, c->stream));
    if (range) {
      LaunchScope ls(c, K_SEARCH_SCAN);
      search_multi_kernel<2><<<blocks, SM_THREADS, smem, c->stream>>>(a);
    }
    {
      LaunchScope ls(c, K_SEARCH_FIRST);
      search_first_rows_kernel<<<qb, 256, 0, c->stream>>>(tot, d_q2u, nq, qcnt);
    }
    {
      LaunchScope ls(c, K_SCAN_COUNTS);
      scan_counts_kernel<<<(unsigned)std::min<uint64_t>(scan_tiles, 1024), 256, 0, c->stream>>>(
          qcnt, nq, c->out_off.as<unsigned long long>(), &c->d_counters->aux[2], c->status.as<uint64_t>(), c->d_counters);
    }
    if (!sizes_only) {
      DMG_TRY(c->out_ids.reserve(std::max<uint64_t>(dev_cap, 1) * 4));
      LaunchScope ls(c, K_SEARCH_EMIT);
      search_first_emit_kernel<<<qb, 256, 0, c->stream>>>(tot, d_q2u, nq, a.out_off, c->out_ids.as<uint32_t>(), dev_cap);
    }
  } else {
    DMG_TRY(c->tmp1.reserve((size_t)tiles * U * 4 + 256));
    a.cnt = c->tmp1.as<uint32_t>();
    DMG_CUDA(cudaMemsetAsync(a.cnt, 0, (size_t)tiles * U * 4, c->stream));
    if (range) {
      LaunchScope ls(c, K_SEARCH_SCAN);
      search_multi_kernel<0><<<blocks, SM_THREADS, smem, c->stream>>>(a);
    }
    {
      LaunchScope ls(c, K_SEARCH_FIRST);
      search_colscan_kernel<<<(U + 7) / 8, 256, 0, c->stream>>>(a.cnt, a.n_tiles, U, tot);
      search_query_counts_kernel<<<qb, 256, 0, c->stream>>>(tot, d_q2u, nq, qcnt);
    }
    {
      LaunchScope ls(c, K_SCAN_COUNTS);
      scan_counts_kernel<<<(unsigned)std::min<uint64_t>(scan_tiles, 1024), 256, 0, c->stream>>>(
          qcnt, nq, c->out_off.as<unsigned long long>(), &c->d_counters->aux[2], c->status.as<uint64_t>(), c->d_counters);
    }
    if (!sizes_only && range) {
      DMG_TRY(c->out_ids.reserve(std::max<uint64_t>(dev_cap, 1) * 4));
      a.out_ids = c->out_ids.as<uint32_t>();
      search_rows_kernel<<<(U + 255) / 256, 256, 0, c->stream>>>(a.uq_off, a.uq, a.out_off, U, a.prow);
      {
        // logged tiles stream out of the log; tiles with a step that is not in it (a name with more patterns
        // than a list holds, log full) are matched again -- that kernel leaves at once for the others
        LaunchScope ls(c, K_SEARCH_EMIT, true, 2);
        search_emit_log_kernel<<<(unsigned)e->sm_count * 8, 256, 0, c->stream>>>(a);
        search_multi_kernel<1><<<blocks, SM_THREADS, smem, c->stream>>>(a);
      }
      if (n_dup) {
        LaunchScope ls(c, K_SEARCH_REPLICATE);
        search_replicate_kernel<<<dim3(n_dup, 32), 256, 0, c->stream>>>(d_dup_q, d_dup_first, a.out_off, a.out_ids, dev_cap);
      }
    }
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  return DMG_OK;
}

}  // namespace dmg
