// Traversal kernels: adjacency rows (T0/T1), batched multi-source k-hop neighbour resolution
// (T2/T3/T4 of SURVEY.md 8a).  Integer-only, HBM/L2-bound work: no tensor cores.
//
// Layout in HBM: forward CSR (out_ptr/out_idx, rows in addDependency order), reverse CSR
// (in_ptr/in_idx, rows in ascending source id), project range table part_off.  Queries and
// results are CSR-shaped (q_off/q_src -> out_off/out_ids/out_hop).
//
// Work items are handed out through an atomic ticket.  Adjacency rows are placed with the decoupled
// look-back allocator of common.cuh in one pass; k-hop results take three steps -- traverse into a
// compact per-query form (bit rows, hop-byte staging rows, or cursor-ordered rows), scan the row sizes
// into out_off, then emit / gather the packed rows in query order.
#include <algorithm>
#include <cstdlib>
#include <atomic>
#include <mutex>

#include "common.cuh"

namespace dmg {

struct KhopArgs {
  const uint32_t* out_ptr;
  const uint32_t* out_idx;
  const uint32_t* in_ptr;
  const uint32_t* in_idx;
  const uint32_t* und_ptr;   // resolve() rows (out row ++ in row)
  const uint32_t* und_idx;
  const uint32_t* part_off;
  uint32_t V, n_parts;
  int mode, k, exclude;
  unsigned long long nq;
  const uint32_t* q_off;
  const uint32_t* q_src;
  unsigned long long* out_off;
  uint32_t* out_ids;
  uint8_t* out_hop;
  unsigned long long cap;
  uint64_t* status;
  DevCounters* ctr;
  uint32_t win_cap;   // largest project, rounded up to a multiple of 128 vertices
  // khop_small only: query chunk [q_begin, q_end), staging rows, row sizes, (lo, wn) per query
  unsigned long long q_begin, q_end;
  uint8_t* stage;
  uint32_t* counts;
  uint32_t* qwin;
  // khop_global only: rows are written where an atomic cursor puts them (tmp_*), then gathered
  unsigned long long* tmp_start;
  uint32_t* tmp_ids;
  uint8_t* tmp_hop;
};

__device__ __forceinline__ void find_window(const KhopArgs& a, uint32_t qb, uint32_t qe, uint32_t* lo, uint32_t* hi) {
  // The query's graph is the project of its first valid source (CCS:1426-1434, 1558).
  *lo = 0;
  *hi = 0;
  for (uint32_t i = qb; i < qe; ++i) {
    const uint32_t s = a.q_src[i];
    if (s < a.V) {
      uint32_t l = 0, r = a.n_parts;
      while (r - l > 1) {
        const uint32_t m = (l + r) >> 1;
        if (a.part_off[m] <= s) l = m; else r = m;
      }
      *lo = a.part_off[l];
      *hi = a.part_off[l + 1];
      return;
    }
  }
}

// (lo, wn) of every query of [q_begin, q_end), one thread per query, ahead of khop_small: the
// bisection over part_off is a chain of dependent loads that would otherwise sit at the head of
// every query with 127 threads of its CTA waiting for it.
__global__ void __launch_bounds__(256) khop_windows_kernel(const KhopArgs a) {
  const unsigned long long q = a.q_begin + blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (q >= a.q_end) return;
  uint32_t lo, hi;
  find_window(a, a.q_off[q], a.q_off[q + 1], &lo, &hi);
  reinterpret_cast<uint2*>(a.qwin)[q] = make_uint2(lo, hi - lo);
}
static inline void launch_windows(const KhopArgs& a, cudaStream_t stream) {
  const unsigned long long n = a.q_end - a.q_begin;
  if (n) khop_windows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(a);
}

}  // namespace dmg
#include "khop_bits.cuh"
namespace dmg {

// ---------------------------------------------------------------------------------------------
// khop_small: one CTA per query, everything about the query lives in shared memory.
// Used when every project has <= 8192 vertices (BASELINE config 2: 2000-vertex projects).
//   hop8[WC]      u8   0 = unseen, 255 = excluded source, else hop; the only per-vertex state
//                      (a level's frontier is "hop == level - 1", found again by a SWAR sweep)
//   srcb[WC/32]   u32  the distinct sources (level-0 frontier membership for bottom-up level 1)
//   list[WC]      u16  ids relative to the project base: the frontier of a top-down level or the
//                      unseen vertices of a bottom-up level (one shared atomic per warp sweep step)
// Level step, chosen per level (direction-optimising BFS; the result is the same set):
//   top-down   the frontier's rows are flattened: a block scan of the row lengths gives every
//              adjacency entry a position, each warp takes a contiguous quarter of the positions
//              and walks it 64 at a time with the current row's (end, offset) record in
//              registers, so lanes stay busy whatever the degree distribution
//   bottom-up  when the frontier's rows hold many more entries than there are unseen vertices
//              (m_f > BU_ALPHA * n_unseen, BU_ALPHA = 12) each unseen vertex looks for ONE
//              frontier member in its reverse row instead: 8 lanes per vertex, 32-byte chunks,
//              stop at the first hit (one lane per vertex when most vertices are frontier)
// The finished hop8 array is copied to a staging row in global memory; khop_emit_kernel
// compacts the staging rows into the packed, id-ordered output after a scan of the row sizes.
// What bounds it (ncu, config 2): ~80% issue utilisation at 16 CTAs/SM with the 25 MB CSR in L2;
// the kernel answers to instruction count AND to the length of the per-query dependency chain
// (barriers, dependent loads) - folding the frontier-list build into the degree pass removed
// instructions but added a barrier to the chain and ran 10% slower; building the frontier list
// and the unseen list in one sweep (one array filled from both ends) saved a sweep and a barrier
// but appends ~1,800 unseen vertices per level that a top-down level never reads: also 10% slower.
// ---------------------------------------------------------------------------------------------
// tuning knobs of khop_small_kernel (compile time; variants are built with -D for experiments)
#ifndef DMG_KS_THREADS
#define DMG_KS_THREADS 128
#endif
#ifndef DMG_KS_MINB
#define DMG_KS_MINB 16
#endif
constexpr int KS_THREADS = DMG_KS_THREADS;   // threads per query CTA
#ifndef DMG_BU_ALPHA
#define DMG_BU_ALPHA 12
#endif
constexpr uint32_t BU_ALPHA = DMG_BU_ALPHA;

// SWAR helpers on a word of four hop bytes: bit 7 of a byte is set in the result where ...
__device__ __forceinline__ uint32_t hop_bytes_eq(uint32_t x, uint32_t pattern4) {   // ... the byte equals pattern4's
  const uint32_t y = x ^ pattern4;
  return ~(((y & 0x7f7f7f7fu) + 0x7f7f7f7fu) | y) & 0x80808080u;
}
__device__ __forceinline__ uint32_t hop_bytes_member(uint32_t x) {   // ... the byte is neither 0 nor 255
  const uint32_t y = ~x;
  const uint32_t nz = ((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x;
  const uint32_t nf = ((y & 0x7f7f7f7fu) + 0x7f7f7f7fu) | y;
  return nz & nf & 0x80808080u;
}
// Whole warp: appends the vertices (4w + byte) whose byte is marked in m to list[*n ...]; one
// shared atomic per warp call, the order inside the list is immaterial (it is a set).
__device__ __forceinline__ void append_marked(uint16_t* list, uint32_t* n, uint32_t w, uint32_t m, int lane) {
  const uint32_t c = __popc(m);
  const uint32_t inc = warp_incl_scan_u32(c);
  const uint32_t tot = __shfl_sync(FULL, inc, 31);
  if (tot == 0) return;
  uint32_t pos = 0;
  if (lane == 31) pos = atomicAdd(n, tot);
  pos = __shfl_sync(FULL, pos, 31) + inc - c;
  while (m) {
    list[pos++] = (uint16_t)(4 * w + ((__ffs(m) - 1) >> 3));
    m &= m - 1;
  }
}
// keeps the bytes of word w that are vertices of a window of wn vertices (w < ceil(wn / 4))
__device__ __forceinline__ uint32_t hop_bytes_inside(uint32_t w, uint32_t wn) {
  const uint32_t n = min(4u, wn - 4u * w);
  return 0x80808080u >> (8u * (4u - n));
}

template <int THREADS, int WC>
__global__ void __launch_bounds__(THREADS, DMG_KS_MINB) khop_small_kernel(const KhopArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint8_t* const hop8 = smem;                                             // WC bytes
  uint32_t* const hop32 = reinterpret_cast<uint32_t*>(smem);
  uint32_t* const srcb = reinterpret_cast<uint32_t*>(smem + WC);          // WC/8 bytes
  uint16_t* const list = reinterpret_cast<uint16_t*>(smem + WC + WC / 8); // WC u16: frontier or unseen list
  __shared__ unsigned long long s_q;
  __shared__ uint32_t s_n, s_cnt;
  __shared__ uint32_t s_wsum[THREADS / 32 + 1];
  __shared__ uint32_t s_red[3][THREADS / 32];
  __shared__ uint2 s_row[THREADS];   // flattened frontier rows of a chunk: (end position, idx offset - start position)
  constexpr int NW = THREADS / 32, NG = THREADS / 8;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int grp = tid >> 3, gl = tid & 7;
  const unsigned gmask = 0xFFu << (8 * (grp & 3));
  // forward rows (top-down) and reverse rows (bottom-up) for the step function of this mode
  const uint32_t* __restrict__ fptr = a.mode == DMG_BOTH ? a.und_ptr : (a.mode == DMG_OUT ? a.out_ptr : a.in_ptr);
  const uint32_t* __restrict__ fidx = a.mode == DMG_BOTH ? a.und_idx : (a.mode == DMG_OUT ? a.out_idx : a.in_idx);
  const uint32_t* __restrict__ rptr = a.mode == DMG_BOTH ? a.und_ptr : (a.mode == DMG_OUT ? a.in_ptr : a.out_ptr);
  const uint32_t* __restrict__ ridx = a.mode == DMG_BOTH ? a.und_idx : (a.mode == DMG_OUT ? a.in_idx : a.out_idx);
  // statistics live in shared memory (thread 0 / one atomic per warp and query): registers are
  // what limits this kernel to 16 CTAs per SM
  __shared__ unsigned long long s_te, s_rows, s_reads;
  if (tid == 0) s_te = s_rows = s_reads = 0;
  // The ticket of the NEXT query is claimed while the current one is processed.  Tickets are
  // the low word of ctr->ticket (zeroed per launch; a launch covers far fewer than 2^32 queries).
  uint32_t next = 0;
  if (tid == 0) next = atomicAdd(reinterpret_cast<unsigned int*>(&a.ctr->ticket), 1u);

  while (true) {
    if (tid == 0) {
      s_q = a.q_begin + next;
      s_n = 0;
      s_cnt = 0;
    }
    __syncthreads();
    const unsigned long long q = s_q;
    if (q >= a.q_end) break;
    if (tid == 0) next = atomicAdd(reinterpret_cast<unsigned int*>(&a.ctr->ticket), 1u);
    uint32_t reads = 0;
    const uint2 win = reinterpret_cast<const uint2*>(a.qwin)[q];   // written by khop_windows_kernel
    const uint32_t qb = a.q_off[q], qe = a.q_off[q + 1];
    // hop bytes and source bitmap are contiguous: clear all of both while those loads are in flight
    for (uint32_t i = tid; i < (WC + WC / 8) / 16; i += THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    const uint32_t lo = win.x, wn = win.y, hi = lo + wn;
    const uint32_t nw4 = (wn + 3) >> 2;
    __syncthreads();
    // L0: distinct sources inside the project -> list
    for (uint32_t i = qb + tid; i < qe; i += THREADS) {
      const uint32_t s = a.q_src[i];
      if (s >= lo && s < hi) {
        const uint32_t v = s - lo, bit = 1u << (v & 31);
        if (!(atomicOr(&srcb[v >> 5], bit) & bit)) {
          list[atomicAdd(&s_n, 1u)] = (uint16_t)v;
          if (a.exclude) hop8[v] = 255;
        }
      }
    }
    __syncthreads();
    const uint32_t ns = s_n;

    for (int lvl = 1; lvl <= a.k; ++lvl) {
      // ---- pass A: frontier size nf, its adjacency entries m_f (TE of SURVEY.md 8d), unseen count
      const uint32_t want = (uint32_t)(lvl - 1), want4 = want * 0x01010101u;
      uint32_t mf = 0, nf = 0, nun = 0;
      if (lvl == 1) {
        for (uint32_t i = tid; i < ns; i += THREADS) {
          const uint32_t u = lo + list[i];
          mf += fptr[u + 1] - fptr[u];
        }
      } else {
        for (uint32_t w = tid; w < nw4; w += THREADS) {
          const uint32_t x = hop32[w];
          uint32_t fm = hop_bytes_eq(x, want4);
          nf += __popc(fm);
          nun += __popc(hop_bytes_eq(x, 0u) & hop_bytes_inside(w, wn));
          while (fm) {
            const uint32_t u = lo + 4 * w + ((__ffs(fm) - 1) >> 3);
            fm &= fm - 1;
            mf += fptr[u + 1] - fptr[u];
          }
        }
      }
      mf = warp_sum_u32(mf);
      nf = warp_sum_u32(nf);
      nun = warp_sum_u32(nun);
      if (lane == 0) {
        s_red[0][warp] = mf;
        s_red[1][warp] = nf;
        s_red[2][warp] = nun;
      }
      __syncthreads();
      mf = nf = nun = 0;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        mf += s_red[0][w];
        nf += s_red[1][w];
        nun += s_red[2][w];
      }
      if (lvl == 1) {
        nf = ns;
        nun = wn - (a.exclude ? ns : 0u);
      }
      if (tid == 0) s_n = 0;
      __syncthreads();   // s_red may be rewritten by the next level; s_n reset is visible
      if (nf == 0) break;
      if (tid == 0) {
        s_te += mf;
        s_rows += nf;
      }
      const bool bottom_up = mf > BU_ALPHA * nun;

      if (!bottom_up) {
        if (lvl > 1) {   // frontier = vertices found by the previous level -> list
          for (uint32_t w0 = 0; w0 < nw4; w0 += THREADS) {
            const uint32_t w = w0 + tid;
            append_marked(list, &s_n, w, (w < nw4) ? hop_bytes_eq(hop32[w], want4) : 0u, lane);
          }
          __syncthreads();
        }
        for (uint32_t c0 = 0; c0 < nf; c0 += THREADS) {
          const uint32_t i = c0 + tid;
          uint32_t b = 0, n = 0;
          if (i < nf) {
            const uint32_t u = lo + list[i];
            b = fptr[u];
            n = fptr[u + 1] - b;
          }
          uint32_t T;
          const uint32_t pre = block_excl_scan<THREADS>(n, s_wsum, &T);
          s_row[tid] = make_uint2(pre + n, b - pre);
          __syncthreads();
          const uint32_t Tq = (((T + NW - 1) / NW) + 31u) & ~31u;
          const uint32_t pbeg = warp * Tq, pend = min(T, pbeg + Tq);
          if (pbeg + lane < pend) {
            uint32_t r;   // first row whose end lies beyond pbeg
            {
              uint32_t l = 0, h = THREADS;
              while (l < h) {
                const uint32_t m = (l + h) >> 1;
                if (s_row[m].x <= pbeg) l = m + 1; else h = m;
              }
              r = l;
            }
            reads += (pend - pbeg - lane + 31u) >> 5;
            uint2 row = s_row[r];   // stays in registers while the lane's positions stay inside the row
            for (uint32_t p = pbeg + lane; p < pend; p += 64) {   // two entries in flight per lane
              while (row.x <= p) row = s_row[++r];
              const uint32_t i0 = fidx[row.y + p];
              const uint32_t p1 = p + 32;
              if (p1 < pend) {
                while (row.x <= p1) row = s_row[++r];
                const uint32_t v1 = fidx[row.y + p1] - lo;
                if (hop8[v1] == 0) hop8[v1] = (uint8_t)lvl;
              }
              const uint32_t v0 = i0 - lo;
              if (hop8[v0] == 0) hop8[v0] = (uint8_t)lvl;   // same value from every racer
            }
          }
          __syncthreads();   // s_row is rewritten by the next chunk
        }
      } else {
        // unseen vertices -> list
        for (uint32_t w0 = 0; w0 < nw4; w0 += THREADS) {
          const uint32_t w = w0 + tid;
          append_marked(list, &s_n, w, (w < nw4) ? (hop_bytes_eq(hop32[w], 0u) & hop_bytes_inside(w, wn)) : 0u, lane);
        }
        __syncthreads();
        const uint32_t nu = s_n;
        if (2 * nf > wn) {
          // dense frontier: the first entry of a row is almost surely a member -> one lane per vertex
          for (uint32_t i = tid; i < nu; i += THREADS) {
            const uint32_t v = list[i];
            const uint32_t b = rptr[lo + v], e = rptr[lo + v + 1];
            for (uint32_t j = b; j < e; ++j) {
              const uint32_t u = ridx[j] - lo;
              ++reads;
              const bool hit = (lvl == 1) ? ((srcb[u >> 5] >> (u & 31)) & 1u) : (hop8[u] == want);
              if (hit) {
                hop8[v] = (uint8_t)lvl;
                break;
              }
            }
          }
        } else {
          // 8 lanes per unseen vertex, 32-byte chunks of its reverse row, stop at the first member
          uint32_t i = grp, v = 0, b = 0, n = 0, j0 = 0;
          bool active = i < nu;
          if (active) {
            v = list[i];
            b = rptr[lo + v];
            n = rptr[lo + v + 1] - b;
          }
          while (__any_sync(FULL, active)) {
            bool hit = false;
            if (active && j0 + gl < n) {
              const uint32_t u = ridx[b + j0 + gl] - lo;
              ++reads;
              hit = (lvl == 1) ? ((srcb[u >> 5] >> (u & 31)) & 1u) : (hop8[u] == want);
            }
            const bool ghit = (__ballot_sync(FULL, hit) & gmask) != 0;
            if (active) {
              if (ghit && gl == 0) hop8[v] = (uint8_t)lvl;
              if (ghit || j0 + 8 >= n) {
                i += NG;
                active = i < nu;
                j0 = 0;
                if (active) {
                  v = list[i];
                  b = rptr[lo + v];
                  n = rptr[lo + v + 1] - b;
                }
              } else {
                j0 += 8;
              }
            }
          }
        }
      }
      __syncthreads();
    }

    // stage the finished hop array and the row size; khop_emit_kernel writes the packed row
    uint32_t* st = reinterpret_cast<uint32_t*>(a.stage + (size_t)(q - a.q_begin) * a.win_cap);
    uint32_t cnt = 0;
    for (uint32_t i = tid; i < nw4; i += THREADS) {
      const uint32_t x = hop32[i];
      st[i] = x;
      cnt += __popc(hop_bytes_member(x));
    }
    cnt = warp_sum_u32(cnt);
    reads = warp_sum_u32(reads);
    if (lane == 0) {
      if (cnt) atomicAdd(&s_cnt, cnt);
      if (reads) atomicAdd(&s_reads, (unsigned long long)reads);
    }
    __syncthreads();
    if (tid == 0) a.counts[q] = s_cnt;
  }
  if (tid == 0) {   // the barrier that ended the loop follows every update of the three tallies
    if (s_rows) {
      atomicAdd(&a.ctr->edges, s_te);
      atomicAdd(&a.ctr->rows, s_rows);
    }
    if (s_reads) atomicAdd(&a.ctr->aux[0], s_reads);
  }
}

// picks the instantiation whose compile-time window covers the largest project
template <class F>
static inline int32_t with_small_kernel(uint32_t win_cap, F f) {
  if (win_cap <= 512) return f(khop_small_kernel<KS_THREADS, 512>, 512);
  if (win_cap <= 2048) return f(khop_small_kernel<KS_THREADS, 2048>, 2048);
  if (win_cap <= 4096) return f(khop_small_kernel<KS_THREADS, 4096>, 4096);   // 12.5 KB: still 16 CTAs/SM
  return f(khop_small_kernel<KS_THREADS, 8192>, 8192);
}
static inline size_t small_kernel_smem(uint32_t wc) { return (size_t)wc + wc / 8 + 2 * (size_t)wc; }

// One warp per query: staging row (hop bytes in id order) -> packed (id, hop) row.  A warp step
// takes 512 vertices as four coalesced word loads per lane; each word is then dealt out 32
// vertices at a time (one shuffle puts vertex 32r+lane in lane `lane`), a ballot ranks the
// members, and the ids / hops go straight to the packed row as contiguous warp-wide stores.
// The kernel is bound by the write-only HBM rate (profiles/write_bw_probe.py: a plain 600 MB
// fill runs at 3.7 TB/s on this part, against 5.9 TB/s for a copy), not by its instructions:
// a shared-memory compaction, this version and a software-pipelined one all take 0.24 ms on
// config 2.  Running these writes underneath the traversal (one persistent launch with
// interleaved traverse / emit tickets and a per-chunk scan) was measured too: 1.14-1.23 ms per
// step against 1.12 ms for the three launches, because the traversal is latency-bound and
// loses the warp slots the emit work holds.
constexpr int EM_THREADS = 256;
constexpr int EM_BLOCKS_PER_SM = 8;   // resident CTAs
constexpr int EM_GRID_PER_SM = 64;    // grid cap: several small waves balance better than one (0.240 vs 0.256 ms)
__global__ void __launch_bounds__(EM_THREADS, EM_BLOCKS_PER_SM) khop_emit_kernel(const uint8_t* __restrict__ stage, uint32_t WC,
                                                               unsigned long long q_begin, unsigned long long q_end,
                                                               const uint32_t* __restrict__ qwin,
                                                               const unsigned long long* __restrict__ out_off,
                                                               uint32_t* out_ids, uint8_t* out_hop,
                                                               unsigned long long cap) {
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
  const uint32_t src_lane = (uint32_t)lane >> 2, sh = 8u * ((uint32_t)lane & 3u);
  const unsigned lt = lanemask_lt();
  for (unsigned long long q = q_begin + warp; q < q_end; q += nwarps) {
    const uint32_t lo = qwin[2 * q], wn = qwin[2 * q + 1], nw4 = (wn + 3) >> 2;
    const uint32_t* __restrict__ st = reinterpret_cast<const uint32_t*>(stage + (size_t)(q - q_begin) * WC);
    const unsigned long long row = out_off[q];
    if (row >= cap) continue;
    const uint32_t room = (uint32_t)min(cap - row, 0xFFFFFFFFull);   // entries of this row that fit
    uint32_t* __restrict__ ids = out_ids + row;
    uint8_t* __restrict__ hops = out_hop + row;
    uint32_t run = 0;
    for (uint32_t w0 = 0; w0 < nw4; w0 += 128) {
      uint32_t x[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {   // words past the row are stale staging memory
        const uint32_t w = w0 + 32 * k + lane;
        x[k] = (w < nw4) ? st[w] : 0u;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (w0 + 32 * k >= nw4) break;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const uint32_t h = (__shfl_sync(FULL, x[k], 8 * r + src_lane) >> sh) & 255u;
          const bool member = (h - 1u) < 254u;   // 0 = not reached, 255 = excluded source
          const unsigned m = __ballot_sync(FULL, member);
          const uint32_t o = run + __popc(m & lt);
          if (member && o < room) {
            ids[o] = lo + 4 * (w0 + 32 * k) + 32 * r + lane;
            hops[o] = (uint8_t)h;
          }
          run += __popc(m);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// khop_global: one CTA per query for projects too large for shared memory.  Per-CTA scratch in
// global memory (L2-resident for the common sizes): claim bitmap over the project, a queue that
// holds sources + every discovered vertex in discovery order (level segments), and a hop byte
// array used only when the row is too long to sort in shared memory.  The bitmap is reset
// sparsely (by walking the queue), so the cost of a query follows its result, not V.
// ---------------------------------------------------------------------------------------------
#ifndef DMG_KG_THREADS
#define DMG_KG_THREADS 512
#endif
constexpr int KG_THREADS = DMG_KG_THREADS;   // threads per query CTA (tuning knob, like DMG_KS_*)
constexpr int KG_SORT_CAP = 4096;    // rows up to this length are ordered in shared memory
constexpr int KG_BUCKETS = 1024;
constexpr uint32_t KG_BUCKET_MAX = 48;   // longest bucket a single thread insertion-sorts

__global__ void __launch_bounds__(KG_THREADS) khop_global_kernel(const KhopArgs a, uint32_t* sc_seen, uint32_t* sc_rank,
                                                                 uint32_t* sc_queue, size_t seen_words, size_t qcap,
                                                                 const uint32_t* qlist, const uint32_t* qlist_n) {
  constexpr int THREADS = KG_THREADS, NW = THREADS / 32;
  extern __shared__ __align__(16) unsigned char kg_smem[];
  uint32_t* s_ids = reinterpret_cast<uint32_t*>(kg_smem);                    // [KG_SORT_CAP]
  uint32_t* s_hist = s_ids + KG_SORT_CAP;                                    // [KG_BUCKETS + 1]
  uint32_t* s_cur = s_hist + KG_BUCKETS + 8;                                 // [KG_BUCKETS]
  uint8_t* s_hopv = reinterpret_cast<uint8_t*>(s_cur + KG_BUCKETS);          // [KG_SORT_CAP]
  __shared__ unsigned long long s_q, s_base;
  __shared__ uint32_t s_lo, s_hi, s_tail, s_flag;
  __shared__ uint32_t s_seg[DMG_MAX_HOPS + 2];
  __shared__ uint32_t s_wsum[NW + 1];
  __shared__ uint2 s_row[THREADS];   // (end position, idx offset - start position) per flattened row
  uint32_t* seen = sc_seen + (size_t)blockIdx.x * seen_words;
  uint32_t* rankw = sc_rank + (size_t)blockIdx.x * seen_words;   // per-word exclusive popcounts of a long row
  uint32_t* queue = sc_queue + (size_t)blockIdx.x * qcap;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t* __restrict__ fptr = a.mode == DMG_BOTH ? a.und_ptr : (a.mode == DMG_OUT ? a.out_ptr : a.in_ptr);
  const uint32_t* __restrict__ fidx = a.mode == DMG_BOTH ? a.und_idx : (a.mode == DMG_OUT ? a.out_idx : a.in_idx);
  unsigned long long te = 0, rows = 0;
  // list mode (the k = 1 queries khop1_warp_kernel handed on): tickets index the list
  const unsigned long long n_run = qlist ? (unsigned long long)*qlist_n : a.nq;

  while (true) {
    if (tid == 0) s_q = atomicAdd(&a.ctr->ticket, 1ull);
    __syncthreads();
    const unsigned long long ticket = s_q;
    if (ticket >= n_run) break;
    const unsigned long long q = qlist ? qlist[ticket] : ticket;
    const uint32_t qb = a.q_off[q], qe = a.q_off[q + 1];
    if (tid == 0) {
      uint32_t lo, hi;
      find_window(a, qb, qe, &lo, &hi);
      s_lo = lo;
      s_hi = hi;
      s_tail = 0;
    }
    __syncthreads();
    const uint32_t lo = s_lo, hi = s_hi, wn = hi - lo;
    for (uint32_t i = qb + tid; i < qe; i += THREADS) {
      const uint32_t s = a.q_src[i];
      if (s >= lo && s < hi) {
        const uint32_t v = s - lo, bit = 1u << (v & 31);
        if (!(atomicOr(&seen[v >> 5], bit) & bit)) queue[atomicAdd(&s_tail, 1u)] = v;
      }
    }
    __syncthreads();
    const uint32_t ns = s_tail;
    uint32_t head = 0, tail = ns;
    __syncthreads();
    if (!a.exclude) {
      for (uint32_t i = tid; i < ns; i += THREADS) {
        const uint32_t v = queue[i];
        atomicAnd(&seen[v >> 5], ~(1u << (v & 31)));
      }
      __threadfence_block();
      __syncthreads();
    }
    if (tid == 0) s_seg[0] = ns;
    int nlev = 0;
    for (int lvl = 1; lvl <= a.k && head < tail; ++lvl) {
      // the frontier's rows are flattened chunk by chunk (see khop_small_kernel)
      for (uint32_t c0 = head; c0 < tail; c0 += THREADS) {
        const uint32_t i = c0 + tid;
        uint32_t b = 0, n = 0;
        if (i < tail) {
          const uint32_t u = lo + queue[i];
          b = fptr[u];
          n = fptr[u + 1] - b;
        }
        uint32_t T;
        const uint32_t pre = block_excl_scan<THREADS>(n, s_wsum, &T);
        s_row[tid] = make_uint2(pre + n, b - pre);
        if (tid == 0) {
          te += T;
          rows += min(tail - c0, (uint32_t)THREADS);
        }
        __syncthreads();
        const uint32_t Tq = (((T + NW - 1) / NW) + 31u) & ~31u;
        const uint32_t pbeg = warp * Tq, pend = min(T, pbeg + Tq);
        if (pbeg < pend) {
          uint32_t r;
          {
            uint32_t l = 0, h = THREADS;
            while (l < h) {
              const uint32_t m = (l + h) >> 1;
              if (s_row[m].x <= pbeg) l = m + 1; else h = m;
            }
            r = l;
          }
          uint2 row = s_row[r];   // stays in registers while the lane's positions stay inside the row
          for (uint32_t p0 = pbeg; p0 < pend; p0 += 32) {
            const uint32_t p = p0 + lane;
            bool isnew = false;
            uint32_t v = 0;
            if (p < pend) {
              while (row.x <= p) row = s_row[++r];
              v = fidx[row.y + p] - lo;
              const uint32_t bit = 1u << (v & 31);
              if (!(__ldcg(&seen[v >> 5]) & bit)) isnew = !(atomicOr(&seen[v >> 5], bit) & bit);
            }
            const unsigned m = __ballot_sync(FULL, isnew);
            if (m) {
              uint32_t pos = 0;
              if (lane == 0) pos = atomicAdd(&s_tail, (uint32_t)__popc(m));
              pos = __shfl_sync(FULL, pos, 0);
              if (isnew) queue[pos + __popc(m & lanemask_lt())] = v;
            }
          }
        }
        __syncthreads();
      }
      head = tail;
      tail = s_tail;
      nlev = lvl;
      if (tid == 0) s_seg[lvl] = tail;   // level lvl is queue[s_seg[lvl-1], s_seg[lvl])
      __syncthreads();
    }
    const uint32_t total = tail - ns;
    if (tid == 0) {
      // rows land wherever the cursor is (no waiting for earlier queries); khop_gather_kernel puts
      // them in query order once scan_counts_kernel has turned the sizes into out_off
      const unsigned long long at = atomicAdd(&a.ctr->aux[3], (unsigned long long)total);
      s_base = at;
      a.tmp_start[q] = at;
      a.counts[q] = total;
    }
    if (tid == 0) s_flag = 0;
    for (uint32_t i = tid; i <= KG_BUCKETS; i += THREADS) s_hist[i] = 0;
    __syncthreads();
    const unsigned long long base = s_base;
    bool scan_path = total > (uint32_t)KG_SORT_CAP;
    if (total && !scan_path) {
      // short row: one MSD radix pass into <= 1024 id buckets, then each bucket is a handful of
      // ids that one thread orders in place
      uint32_t sh = 0;   // the largest id, wn - 1, must land in bucket KG_BUCKETS - 1 at most
      while (((wn - 1) >> sh) >= (uint32_t)KG_BUCKETS) ++sh;
      for (uint32_t i = tid; i < total; i += THREADS) atomicAdd(&s_hist[queue[ns + i] >> sh], 1u);
      __syncthreads();
      {
        constexpr int PER = KG_BUCKETS / THREADS;
        uint32_t x[PER], sum = 0;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
          x[j] = s_hist[tid * PER + j];
          sum += x[j];
        }
        uint32_t tot;
        uint32_t off = block_excl_scan<THREADS>(sum, s_wsum, &tot);
#pragma unroll
        for (int j = 0; j < PER; ++j) {
          s_hist[tid * PER + j] = off;
          s_cur[tid * PER + j] = off;
          if (x[j] > KG_BUCKET_MAX) s_flag = 1;
          off += x[j];
        }
        if (tid == THREADS - 1) s_hist[KG_BUCKETS] = off;
      }
      __syncthreads();
      for (uint32_t i = tid; i < total; i += THREADS) {
        const uint32_t p = ns + i, v = queue[p];
        int lv = 1;
        while (lv < nlev && p >= s_seg[lv]) ++lv;
        const uint32_t pos = atomicAdd(&s_cur[v >> sh], 1u);
        s_ids[pos] = v;
        s_hopv[pos] = (uint8_t)lv;
      }
      __syncthreads();
      if (!s_flag) {
        for (uint32_t bkt = tid; bkt < (uint32_t)KG_BUCKETS; bkt += THREADS) {
          const uint32_t b0 = s_hist[bkt], b1 = s_hist[bkt + 1];
          for (uint32_t i = b0 + 1; i < b1; ++i) {
            const uint32_t v = s_ids[i];
            const uint8_t h = s_hopv[i];
            uint32_t j = i;
            while (j > b0 && s_ids[j - 1] > v) {
              s_ids[j] = s_ids[j - 1];
              s_hopv[j] = s_hopv[j - 1];
              --j;
            }
            s_ids[j] = v;
            s_hopv[j] = h;
          }
        }
        __syncthreads();
        for (uint32_t i = tid; i < total; i += THREADS) {
          const unsigned long long pos = base + i;
          if (pos < a.cap) {
            a.tmp_ids[pos] = lo + s_ids[i];
            a.tmp_hop[pos] = s_hopv[i];
          }
        }
      } else {
        scan_path = true;   // ids clustered in few buckets: use the bitmap scan instead
      }
    }
    if (total && scan_path) {
      // long row: every discovered vertex is RANKED inside the claim bitmap instead of sorted.
      // Pass 1 turns the bitmap into per-word exclusive popcounts (coalesced, uniform work);
      // pass 2 streams the queue and drops (id, hop) at base + rank(id).  No hop array, no
      // per-bit lookups, one barrier pair per 4096 bitmap words.
      if (a.exclude) {   // excluded sources keep their claim bit during the traversal; they are not results
        for (uint32_t i = tid; i < ns; i += THREADS) {
          const uint32_t v = queue[i];
          atomicAnd(&seen[v >> 5], ~(1u << (v & 31)));
        }
        __threadfence_block();
        __syncthreads();
      }
      const uint32_t nquads = (wn + 127) >> 7;   // the bitmap is padded with zero words to a multiple of 4
      const uint4* seen4 = reinterpret_cast<const uint4*>(seen);
      uint4* rank4 = reinterpret_cast<uint4*>(rankw);
      uint32_t run = 0;
      for (uint32_t g0 = 0; g0 < nquads; g0 += THREADS) {
        const uint32_t gq = g0 + tid;
        uint4 c = make_uint4(0u, 0u, 0u, 0u);
        if (gq < nquads) {
          const uint4 x = __ldcg(seen4 + gq);
          c = make_uint4(__popc(x.x), __popc(x.y), __popc(x.z), __popc(x.w));
        }
        uint32_t tot;
        const uint32_t off = run + block_excl_scan<THREADS>(c.x + c.y + c.z + c.w, s_wsum, &tot);
        if (gq < nquads) rank4[gq] = make_uint4(off, off + c.x, off + c.x + c.y, off + c.x + c.y + c.z);
        run += tot;
      }
      __threadfence_block();
      __syncthreads();
      for (uint32_t i = tid; i < total; i += THREADS) {
        const uint32_t p = ns + i, v = queue[p], w = v >> 5;
        int lv = 1;
        while (lv < nlev && p >= s_seg[lv]) ++lv;
        const unsigned long long pos = base + __ldcg(&rankw[w]) + __popc(__ldcg(&seen[w]) & ((1u << (v & 31)) - 1u));
        if (pos < a.cap) {
          a.tmp_ids[pos] = lo + v;
          a.tmp_hop[pos] = (uint8_t)lv;
        }
      }
    }
    __syncthreads();
    // sparse reset of the claim bitmap
    for (uint32_t i = tid; i < tail; i += THREADS) seen[queue[i] >> 5] = 0;
    __threadfence_block();
    __syncthreads();
  }
  te = warp_sum_u64(te);
  rows = warp_sum_u64(rows);
  if (lane == 0 && rows) {
    atomicAdd(&a.ctr->edges, te);
    atomicAdd(&a.ctr->aux[0], te);   // this kernel is top-down only: read == traversed
    atomicAdd(&a.ctr->rows, rows);
  }
}

// Rows of khop_global_kernel from cursor order into query order.  Work is split by OUTPUT
// position, not by query (rows run from nothing to a whole project): a warp takes 2048 output
// positions, finds the row of the first one by bisection over out_off and walks forward.
constexpr int GA_CHUNK = 2048;
__global__ void __launch_bounds__(256) khop_gather_kernel(unsigned long long nq,
                                                          const unsigned long long* __restrict__ tmp_start,
                                                          const unsigned long long* __restrict__ out_off,
                                                          const uint32_t* __restrict__ tmp_ids,
                                                          const uint8_t* __restrict__ tmp_hop, uint32_t* out_ids,
                                                          uint8_t* out_hop, unsigned long long cap) {
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
  const unsigned long long total = min(out_off[nq], cap);
  for (unsigned long long P = warp * GA_CHUNK; P < total; P += nwarps * GA_CHUNK) {
    const unsigned long long Pend = min(total, P + GA_CHUNK);
    unsigned long long q = 0, hi = nq;   // first row whose end lies beyond P (empty rows are skipped)
    while (q < hi) {
      const unsigned long long mid = (q + hi) >> 1;
      if (out_off[mid + 1] <= P) q = mid + 1; else hi = mid;
    }
    unsigned long long row_end = out_off[q + 1], from = tmp_start[q] - out_off[q];   // u64 wrap-around is intended
    for (unsigned long long p = P + lane; p < Pend; p += 32) {
      while (p >= row_end) {
        ++q;
        row_end = out_off[q + 1];
        from = tmp_start[q] - out_off[q];
      }
      if (p + from < cap) {
        out_ids[p] = tmp_ids[p + from];
        out_hop[p] = tmp_hop[p + from];
      }
    }
  }
}

}  // namespace dmg
#include "khop_large.cuh"
namespace dmg {

// ---------------------------------------------------------------------------------------------
// adjacency_gather: rows of one CSR for a batch of ids (T0 dependencies / T1 dependents).
// One ticket = a tile of THREADS ids; the tile's rows are sized, placed by look-back, then
// copied warp-per-row.
// ---------------------------------------------------------------------------------------------
constexpr int ADJ_THREADS = 256;
constexpr uint32_t ADJ_LONG_ROW = 512;

__global__ void __launch_bounds__(ADJ_THREADS) adjacency_kernel(const uint32_t* __restrict__ ptr,
                                                                const uint32_t* __restrict__ idx, uint32_t V,
                                                                unsigned long long n, const uint32_t* __restrict__ ids,
                                                                unsigned long long* out_off, uint32_t* out_ids,
                                                                unsigned long long cap, uint64_t* status,
                                                                DevCounters* ctr, unsigned long long* long_rows) {
  constexpr int THREADS = ADJ_THREADS, NW = THREADS / 32;
  __shared__ unsigned long long s_t, s_base;
  __shared__ uint32_t s_start[THREADS], s_len[THREADS];
  __shared__ unsigned long long s_off[THREADS];
  __shared__ unsigned long long s_wsum[NW + 1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned long long n_tiles = (n + THREADS - 1) / THREADS;
  unsigned long long te = 0;
  while (true) {
    if (tid == 0) s_t = atomicAdd(&ctr->ticket, 1ull);
    __syncthreads();
    const unsigned long long t = s_t;
    if (t >= n_tiles) break;
    const unsigned long long row = t * THREADS + tid;
    uint32_t start = 0, len = 0;
    if (row < n) {
      const uint32_t id = ids[row];
      if (id < V) {   // unknown id -> empty row (PG:300-302, 317-319)
        start = ptr[id];
        len = ptr[id + 1] - start;
      }
    }
    unsigned long long total;   // 64 bits: a tile of repeated hub ids can exceed 2^32 entries
    const unsigned long long off = block_excl_scan_u64<THREADS>(len, s_wsum, &total);
    s_start[tid] = start;
    s_len[tid] = len;
    s_off[tid] = off;
    if (warp == 0) {
      const unsigned long long excl = lookback_exclusive(status, t, total);
      if (lane == 0) {
        s_base = excl;
        if (t == n_tiles - 1) {
          out_off[n] = excl + total;
          ctr->total = excl + total;
        }
      }
    }
    __syncthreads();
    const unsigned long long base = s_base;
    if (row < n) out_off[row] = base + off;
    te += len;
    for (int r = warp; r < THREADS; r += NW) {
      const uint32_t st = s_start[r], ln = s_len[r];
      const unsigned long long o = base + s_off[r];
      if (ln > ADJ_LONG_ROW) {
        // hub row (Zipf in-degree): left to adjacency_long_kernel, which spreads it over the grid
        if (lane == 0) {
          const unsigned long long slot = atomicAdd(&ctr->aux[1], 1ull);
          long_rows[2 * slot] = ((unsigned long long)st << 32) | ln;
          long_rows[2 * slot + 1] = o;
        }
        continue;
      }
      for (uint32_t j = lane; j < ln; j += 32)
        if (o + j < cap) out_ids[o + j] = idx[st + j];
    }
  }
  te = warp_sum_u64(te);
  if (lane == 0 && te) {
    atomicAdd(&ctr->edges, te);
    atomicAdd(&ctr->aux[0], te);
  }
}

// Rows longer than ADJ_LONG_ROW are copied in chunks of ADJ_CHUNK entries, one warp per chunk, whatever row
// a chunk belongs to: a Zipf hub with 800 k dependents and a few hundred 600-entry rows then cost what
// their entries cost (the previous kernel walked the long rows one after the other with the whole grid on
// each: config 4's `dependents` batch spent most of its 0.26 ms there).
constexpr uint32_t ADJ_CHUNK = 512;
// chunk_off[i] = chunks of the long rows before row i (one CTA: there are at most n rows, usually a handful)
__global__ void __launch_bounds__(1024) adjacency_long_scan_kernel(const unsigned long long* __restrict__ long_rows,
                                                                   DevCounters* ctr, uint32_t* chunk_off) {
  __shared__ uint32_t s_wsum[33];
  const unsigned long long n_long = ctr->aux[1];
  uint32_t run = 0;
  for (unsigned long long i0 = 0; i0 < n_long; i0 += 1024) {
    const unsigned long long i = i0 + threadIdx.x;
    const uint32_t nch = i < n_long ? ((uint32_t)long_rows[2 * i] + ADJ_CHUNK - 1) / ADJ_CHUNK : 0u;
    uint32_t tot;
    const uint32_t off = block_excl_scan<1024>(nch, s_wsum, &tot);
    if (i < n_long) chunk_off[i] = run + off;
    run += tot;
  }
  if (threadIdx.x == 0) ctr->aux[3] = run;   // chunks in all
}
__global__ void __launch_bounds__(256) adjacency_long_kernel(const uint32_t* __restrict__ idx,
                                                             const unsigned long long* __restrict__ long_rows,
                                                             const uint32_t* __restrict__ chunk_off,
                                                             const DevCounters* ctr, uint32_t* out_ids,
                                                             unsigned long long cap) {
  const int lane = threadIdx.x & 31;
  const unsigned long long n_long = ctr->aux[1], n_chunks = ctr->aux[3];
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  const unsigned long long n_warps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
  for (unsigned long long c = warp; c < n_chunks; c += n_warps) {
    unsigned long long r = 0, hi = n_long;   // last row whose first chunk is <= c
    while (hi - r > 1) {
      const unsigned long long mid = (r + hi) >> 1;
      if (chunk_off[mid] <= c) r = mid; else hi = mid;
    }
    const uint32_t st = (uint32_t)(long_rows[2 * r] >> 32), ln = (uint32_t)long_rows[2 * r];
    const unsigned long long o = long_rows[2 * r + 1];
    const uint32_t j0 = (uint32_t)(c - chunk_off[r]) * ADJ_CHUNK, j1 = min(ln, j0 + ADJ_CHUNK);
#pragma unroll 4
    for (uint32_t j = j0 + lane; j < j1; j += 32)
      if (o + j < cap) out_ids[o + j] = idx[st + j];
  }
}

__global__ void entry_flags_kernel(const uint32_t* __restrict__ bits, uint32_t V, unsigned long long n,
                                   const uint32_t* __restrict__ ids, uint8_t* out) {
  const unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t id = ids[i];
    out[i] = (id < V) ? (uint8_t)((bits[id >> 5] >> (id & 31)) & 1u) : 0;   // PG:336-338
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
constexpr uint32_t SMALL_WINDOW_MAX = 8192;


static int32_t reset_counters(CallCtx* c) {
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  return DMG_OK;
}
static int32_t fetch_counters(CallCtx* c) {
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  return DMG_OK;
}

// planes needed for hop codes 1..k (binary-coded): k = 1 -> 1, k <= 3 -> 2, k <= 7 -> 3
static inline int hop_planes(int k) { return k <= 1 ? 1 : (k <= 3 ? 2 : (k <= 7 ? 3 : 8)); }

// Set-word width of the bit-parallel path: 128 queries per group when the sets fit shared memory,
// else 64; 0 = the path does not apply (project too large, too many hops, no ELL copy).
static inline int bit_parallel_w64(const dmg_graph* g, int k) {
  const Engine* e = g->eng;
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  if (!g->has_ell || maxp > MB_MAX_WINDOW || hop_planes(k) > MB_MAX_PLANES) return 0;
  const uint32_t wc = (maxp + 127u) & ~127u;
  static const int forced = [] {
    const char* s = getenv("DMG_MB_W64");   // tuning runs only
    return s ? atoi(s) : 0;
  }();
  for (int w64 = 2; w64 >= 1; --w64) {
    if (forced && w64 != forced) continue;
    if (msbfs_smem(wc, hop_planes(k), k, w64) + 2048 <= e->smem_optin) return w64;
  }
  return 0;
}

template <class F>
static inline int32_t with_planes(int np, int w64, F f) {
  if (w64 == 2) {
    if (np == 1) return f(khop_msbfs_kernel<1, 2>, khop_emit_bits_kernel<1>);
    if (np == 2) return f(khop_msbfs_kernel<2, 2>, khop_emit_bits_kernel<2>);
    return f(khop_msbfs_kernel<3, 2>, khop_emit_bits_kernel<3>);
  }
  if (np == 1) return f(khop_msbfs_kernel<1, 1>, khop_emit_bits_kernel<1>);
  if (np == 2) return f(khop_msbfs_kernel<2, 1>, khop_emit_bits_kernel<2>);
  return f(khop_msbfs_kernel<3, 1>, khop_emit_bits_kernel<3>);
}

// Layout of the grouping scratch of one chunk of queries (all u32 unless noted).
struct MbLayout {
  size_t o_qpart, o_qrank, o_hist, o_gbase, o_ngroups, o_grp_win, o_grp_part, o_grp_q, bytes;
  uint64_t max_groups;
  MbLayout(uint64_t nq, uint32_t n_parts, uint32_t slots) {
    auto pad = [](size_t x) { return (x + 255) & ~size_t(255); };
    max_groups = nq / slots + n_parts + 1;
    o_qpart = 0;
    o_qrank = pad(o_qpart + nq * 4);
    o_hist = pad(o_qrank + nq * 4);
    o_gbase = pad(o_hist + (size_t)n_parts * 4);
    o_ngroups = pad(o_gbase + ((size_t)n_parts + 1) * 4);
    o_grp_win = pad(o_ngroups + 4);
    o_grp_part = pad(o_grp_win + max_groups * 8);
    o_grp_q = pad(o_grp_part + max_groups * 4);
    bytes = pad(o_grp_q + max_groups * slots * 4);
  }
};

// The bit-parallel path (khop_bits.cuh) for one chunk [q0, q1) of the batch: classify + group the
// queries by project, traverse 64 / 128 queries of a project per CTA into bit rows + row sizes.  `bits`
// receives (q1 - q0) rows of np * WB words; counts / qwin are indexed by absolute query.
static int32_t launch_bit_traversal(CallCtx* c, dmg_graph* g, int w64, int mode, int k, int exclude, uint64_t q0,
                                    uint64_t q1, const uint32_t* d_q_off, const uint32_t* d_q_src, uint32_t* bits,
                                    uint32_t* counts, uint32_t* qwin, unsigned char* grp_scratch, DevCounters* ctr,
                                    cudaStream_t stream) {
  Engine* e = g->eng;
  const uint64_t n = q1 - q0;
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  const int np = hop_planes(k);
  const uint32_t slots = 64u * (uint32_t)w64;
  const MbLayout L(n, g->info.n_parts, slots);
  const dmg_graph::Ell& el = g->ell[mode == DMG_BOTH ? 0 : (mode == DMG_OUT ? 1 : 2)];
  MbArgs a;
  a.fptr = mode == DMG_BOTH ? g->d_und_ptr : (mode == DMG_OUT ? g->d_out_ptr : g->d_in_ptr);
  a.ell_proj_slice = el.d_proj_slice;
  a.ell_slice_off = el.d_slice_off;
  a.ell_vrow = el.d_vrow;
  a.ell_data = el.d_data;
  a.slots = slots;
  a.part_off = g->d_part_off;
  a.V = g->info.n_vertices;
  a.n_parts = g->info.n_parts;
  a.k = k;
  a.exclude = exclude ? 1 : 0;
  a.nq = n;
  a.q_off = d_q_off + q0;
  a.q_src = d_q_src;
  a.win_cap = (maxp + 127u) & ~127u;
  a.WB = a.win_cap / 32;
  a.qpart = reinterpret_cast<uint32_t*>(grp_scratch + L.o_qpart);
  a.qrank = reinterpret_cast<uint32_t*>(grp_scratch + L.o_qrank);
  a.qwin = qwin + 2 * q0;
  a.part_hist = reinterpret_cast<uint32_t*>(grp_scratch + L.o_hist);
  a.part_gbase = reinterpret_cast<uint32_t*>(grp_scratch + L.o_gbase);
  a.n_groups = reinterpret_cast<uint32_t*>(grp_scratch + L.o_ngroups);
  a.grp_win = reinterpret_cast<uint2*>(grp_scratch + L.o_grp_win);
  a.grp_part = reinterpret_cast<uint32_t*>(grp_scratch + L.o_grp_part);
  a.grp_q = reinterpret_cast<uint32_t*>(grp_scratch + L.o_grp_q);
  a.bits = bits;
  a.counts = counts + q0;
  a.ctr = ctr;
  DMG_CUDA(cudaMemsetAsync(a.part_hist, 0, (size_t)a.n_parts * 4, stream));
  DMG_CUDA(cudaMemsetAsync(a.grp_q, 0xFF, L.max_groups * slots * 4, stream));
  const size_t smem = msbfs_smem(a.win_cap, np, k, w64);
  int occ = 1;
  DMG_TRY(with_planes(np, w64, [&](auto kern, auto) -> int32_t {
    DMG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, MB_THREADS, smem));
    return DMG_OK;
  }));
  const unsigned qblocks = (unsigned)((n + 255) / 256);
  {
    LaunchScope ls(c, K_KHOP_WINDOWS, true, 3);
    mb_classify_kernel<<<qblocks, 256, 0, stream>>>(a);
    mb_groups_kernel<<<1, 1024, 0, stream>>>(a);
    mb_scatter_kernel<<<qblocks, 256, 0, stream>>>(a);
  }
  const uint64_t grid = std::min<uint64_t>(L.max_groups, (uint64_t)e->sm_count * std::max(occ, 1));
  {
    LaunchScope ls(c, K_KHOP_SMALL);
    with_planes(np, w64, [&](auto kern, auto) -> int32_t {
      kern<<<(unsigned)grid, MB_THREADS, smem, stream>>>(a);
      return DMG_OK;
    });
  }
  return DMG_OK;
}

// Runs the k-hop batch on device buffers on c->stream; leaves counters in c->h_counters after
// the caller synchronises.
int32_t run_khop(CallCtx* c, dmg_graph* g, int mode, int k, int exclude, uint64_t nq,
                        const uint32_t* d_q_off, const uint32_t* d_q_src, uint64_t* d_out_off,
                        uint32_t* d_out_ids, uint8_t* d_out_hop, uint64_t cap) {
  Engine* e = c->eng;
  // scan_counts' look-back words for the whole batch (a staging-limited chunk needs fewer); zeroed
  // here together with the counters, so the first chunk launches without further memsets
  const uint64_t scan_tiles_all = (nq + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
  DMG_TRY(c->status.reserve(scan_tiles_all * sizeof(uint64_t)));
  DMG_CUDA(cudaMemsetAsync(c->status.p, 0, scan_tiles_all * sizeof(uint64_t), c->stream));
  DMG_TRY(reset_counters(c));
  KhopArgs a;
  a.out_ptr = g->d_out_ptr;
  a.out_idx = g->d_out_idx;
  a.in_ptr = g->d_in_ptr;
  a.in_idx = g->d_in_idx;
  a.part_off = g->d_part_off;
  a.V = g->info.n_vertices;
  a.n_parts = g->info.n_parts;
  a.mode = mode;
  a.k = k;
  a.exclude = exclude ? 1 : 0;
  a.nq = nq;
  a.q_off = d_q_off;
  a.q_src = d_q_src;
  a.out_off = reinterpret_cast<unsigned long long*>(d_out_off);
  a.out_ids = d_out_ids;
  a.out_hop = d_out_hop;
  a.cap = cap;
  a.status = c->status.as<uint64_t>();
  a.ctr = c->d_counters;
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  a.win_cap = (maxp + 127u) & ~127u;
  a.und_ptr = g->d_und_ptr;
  a.und_idx = g->d_und_idx;
  a.q_begin = 0;
  a.q_end = nq;
  a.stage = nullptr;
  a.counts = nullptr;
  a.qwin = nullptr;
  if (const int w64 = bit_parallel_w64(g, k)) {
    // 64 / 128 queries of one project per CTA -> bit rows + row sizes; scan; bit rows -> packed rows
    const int np = hop_planes(k);
    const uint32_t WB = a.win_cap / 32;
    const uint64_t row_bytes = (uint64_t)np * WB * 4;
    const uint64_t budget = std::max<uint64_t>(std::min<uint64_t>(e->scratch_budget, 2ull << 30), row_bytes);
    const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(nq, budget / row_bytes));
    const MbLayout L(chunk, g->info.n_parts, 64u * (uint32_t)w64);
    DMG_TRY(c->scratch.reserve(chunk * row_bytes));
    DMG_TRY(c->tmp0.reserve(nq * 4));        // row sizes
    DMG_TRY(c->tmp1.reserve(nq * 8));        // (lo, wn) per query
    DMG_TRY(c->tmp2.reserve(L.bytes));       // grouping scratch
    const uint64_t scan_tiles = (chunk + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
    uint32_t* counts = c->tmp0.as<uint32_t>();
    uint32_t* qwin = c->tmp1.as<uint32_t>();
    unsigned long long* carry = &c->d_counters->aux[2];
    for (uint64_t q0 = 0; q0 < nq; q0 += chunk) {
      const uint64_t q1 = std::min<uint64_t>(nq, q0 + chunk);
      if (q0) {   // later chunks reuse the ticket, the scan ticket and the look-back words
        DMG_CUDA(cudaMemsetAsync(&c->d_counters->ticket, 0, 8, c->stream));
        DMG_CUDA(cudaMemsetAsync(&c->d_counters->aux[1], 0, 8, c->stream));
        DMG_CUDA(cudaMemsetAsync(c->status.p, 0, scan_tiles * sizeof(uint64_t), c->stream));
      }
      DMG_TRY(launch_bit_traversal(c, g, w64, mode, k, exclude, q0, q1, d_q_off, d_q_src, c->scratch.as<uint32_t>(), counts,
                                   qwin, c->tmp2.as<unsigned char>(), c->d_counters, c->stream));
      {
        LaunchScope ls(c, K_SCAN_COUNTS);
        const uint64_t tiles = (q1 - q0 + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
        scan_counts_kernel<<<(unsigned)std::min<uint64_t>(tiles, (uint64_t)e->sm_count * 4), SC_THREADS, 0, c->stream>>>(
            counts + q0, q1 - q0, a.out_off + q0, carry, c->status.as<uint64_t>(), c->d_counters);
      }
      if (a.cap) {
        LaunchScope ls(c, K_KHOP_EMIT);
        // about one query per warp, many small waves (a resident wave whose warps walk several queries with
        // the next query's head loaded ahead was tried: 40 registers, 0.217 vs 0.175 ms)
        const uint64_t blocks = std::min<uint64_t>((q1 - q0 + 7) / 8, (uint64_t)e->sm_count * EM_GRID_PER_SM);
        with_planes(np, w64, [&](auto, auto emit) -> int32_t {
          emit<<<(unsigned)blocks, 256, 0, c->stream>>>(c->scratch.as<uint32_t>(), WB, q1 - q0, qwin + 2 * q0, counts + q0,
                                                        a.out_off + q0, a.out_ids, a.out_hop, a.cap);
          return DMG_OK;
        });
      }
    }
  } else if (maxp <= SMALL_WINDOW_MAX) {
    const uint32_t WC = a.win_cap;
    int occ = 0;
    size_t smem = 0;
    DMG_TRY(with_small_kernel(WC, [&](auto kern, uint32_t wc) -> int32_t {
      smem = small_kernel_smem(wc);
      if (smem > 48 * 1024)
        DMG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, KS_THREADS, smem));
      return DMG_OK;
    }));
    // staging rows (one hop byte per vertex of the project) for a chunk of queries
    const uint64_t stage_budget = std::max<uint64_t>(std::min<uint64_t>(e->scratch_budget, 2ull << 30), WC);
    const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(nq, stage_budget / WC));
    DMG_TRY(c->scratch.reserve(chunk * WC));
    DMG_TRY(c->tmp0.reserve(nq * 4));        // row sizes
    DMG_TRY(c->tmp1.reserve(nq * 8));        // (lo, wn) per query
    const uint64_t scan_tiles = (chunk + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
    a.stage = c->scratch.as<uint8_t>();
    a.counts = c->tmp0.as<uint32_t>();
    a.qwin = c->tmp1.as<uint32_t>();
    unsigned long long* carry = &c->d_counters->aux[2];
    for (uint64_t q0 = 0; q0 < nq; q0 += chunk) {
      const uint64_t q1 = std::min<uint64_t>(nq, q0 + chunk);
      a.q_begin = q0;
      a.q_end = q1;
      if (q0) {   // later chunks reuse the ticket, the scan ticket and the look-back words
        DMG_CUDA(cudaMemsetAsync(&c->d_counters->ticket, 0, 8, c->stream));
        DMG_CUDA(cudaMemsetAsync(&c->d_counters->aux[1], 0, 8, c->stream));
        DMG_CUDA(cudaMemsetAsync(c->status.p, 0, scan_tiles * sizeof(uint64_t), c->stream));
      }
      const uint64_t grid = std::min<uint64_t>(q1 - q0, (uint64_t)e->sm_count * std::max(occ, 1));
      {
        LaunchScope ls(c, K_KHOP_WINDOWS);
        launch_windows(a, c->stream);
      }
      {
        LaunchScope ls(c, K_KHOP_SMALL);
        with_small_kernel(WC, [&](auto kern, uint32_t) -> int32_t {
          kern<<<(unsigned)grid, KS_THREADS, smem, c->stream>>>(a);
          return DMG_OK;
        });
      }
      {
        LaunchScope ls(c, K_SCAN_COUNTS);
        const uint64_t tiles = (q1 - q0 + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
        scan_counts_kernel<<<(unsigned)std::min<uint64_t>(tiles, (uint64_t)e->sm_count * 4), SC_THREADS, 0, c->stream>>>(
            a.counts + q0, q1 - q0, a.out_off + q0, carry, c->status.as<uint64_t>(), c->d_counters);
      }
      {
        LaunchScope ls(c, K_KHOP_EMIT);
        const uint64_t warps = q1 - q0;
        const uint64_t blocks = std::min<uint64_t>((warps + 7) / 8, (uint64_t)e->sm_count * EM_GRID_PER_SM);
        khop_emit_kernel<<<(unsigned)blocks, EM_THREADS, 0, c->stream>>>(a.stage, WC, q0, q1, a.qwin, a.out_off, a.out_ids,
                                                                 a.out_hop, a.cap);
      }
    }
  } else {
    // Projects beyond the shared-memory classes.  k = 1 rows are unions of adjacency rows: one warp per
    // query (khop1_warp_kernel), which hands hub rows / long source lists on to a CTA kernel.  That CTA
    // kernel is khop_large_kernel while a project's claim bitmap fits one SM's shared memory, else
    // khop_global_kernel with its bitmap in global memory.
    const size_t seen_words = (((size_t)maxp + 31) / 32 + 3) & ~size_t(3);
    const size_t qcap = 2 * (size_t)maxp;
    const size_t kl_smem = large_kernel_smem(maxp);
    const bool no_large = getenv("DMG_NO_KHOP_LARGE") != nullptr;   // differential tests: khop_global_kernel instead
    // khop_large_kernel runs one query per SM; it pays off when a query has work for 1024 threads.  Short
    // traversals (expected entries ~ degree^k below a few thousand) are chains of dependent steps: more,
    // smaller CTAs in flight (khop_global_kernel) finish such a batch sooner (config 4, k = 2: 1.2 vs 2.5 ms).
    const double deg = (double)g->info.n_edges * (mode == DMG_BOTH ? 2.0 : 1.0) / std::max<uint32_t>(g->info.n_vertices, 1);
    double expect = 1.0;
    for (int i = 0; i < k && expect < 1e9; ++i) expect *= std::max(deg, 1.0);
    const char* force_large = getenv("DMG_FORCE_KHOP_LARGE");   // differential tests
    const bool listed = k == 1;   // the CTA kernel then works off khop1_warp_kernel's list
    const bool use_large = !no_large && kl_smem + 12 * 1024 <= e->smem_optin && (listed || expect >= 4096.0 || force_large);
    const size_t per_block = use_large ? qcap * 4 : 2 * seen_words * 4 + qcap * 4;   // (claim bitmap, rank words,) queue
    int occ = 0;
    const size_t kg_smem = (size_t)KG_SORT_CAP * 5 + (2 * (size_t)KG_BUCKETS + 8) * 4;
    uint64_t grid;
    if (use_large) {
      DMG_CUDA(cudaFuncSetAttribute(khop_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kl_smem));
      DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, khop_large_kernel, KL_THREADS, kl_smem));
      grid = std::min<uint64_t>(nq, (uint64_t)e->sm_count * std::max(occ, 1));
    } else {
      DMG_CUDA(cudaFuncSetAttribute(khop_global_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kg_smem));
      DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, khop_global_kernel, KG_THREADS, kg_smem));
      grid = std::min<uint64_t>(nq, (uint64_t)e->sm_count * std::max(occ, 1));
      // every CTA hits its claim bitmap and rank words at random: keep those of all resident CTAs
      // inside ~3/4 of L2 (1M-vertex project on B200: 372 CTAs run 7.1 ms where 888 run 11 ms)
      const uint64_t l2_fit = (uint64_t)e->l2_bytes * 3 / 4 / (2 * seen_words * 4);
      grid = std::min<uint64_t>(grid, std::max<uint64_t>(l2_fit, (uint64_t)e->sm_count));
    }
    grid = std::max<uint64_t>(1, std::min<uint64_t>(grid, e->scratch_budget / per_block));
    const size_t defer_bytes = listed ? ((nq * 4 + 255) & ~size_t(255)) + 256 : 0;
    DMG_TRY(c->scratch.reserve(grid * per_block + defer_bytes));
    uint32_t *sc_seen = nullptr, *sc_rank = nullptr, *sc_queue = c->scratch.as<uint32_t>();
    if (!use_large) {
      sc_seen = c->scratch.as<uint32_t>();
      sc_rank = sc_seen + grid * seen_words;
      sc_queue = sc_rank + grid * seen_words;
      DMG_CUDA(cudaMemsetAsync(sc_seen, 0, grid * seen_words * 4, c->stream));
    }
    uint32_t *d_defer = nullptr, *d_ndefer = nullptr;
    if (listed) {
      d_ndefer = reinterpret_cast<uint32_t*>(c->scratch.as<unsigned char>() + grid * per_block);
      d_defer = d_ndefer + 64;
      DMG_CUDA(cudaMemsetAsync(d_ndefer, 0, 4, c->stream));
    }
    // rows are first written in cursor order into temporaries of the caller's capacity
    DMG_TRY(c->tmp0.reserve(nq * 4));
    DMG_TRY(c->tmp1.reserve(nq * 8));
    DMG_TRY(c->tmp2.reserve(std::max<uint64_t>(cap, 1) * 4));
    DMG_TRY(c->tmp3.reserve(std::max<uint64_t>(cap, 1)));
    a.counts = c->tmp0.as<uint32_t>();
    a.tmp_start = c->tmp1.as<unsigned long long>();
    a.tmp_ids = c->tmp2.as<uint32_t>();
    a.tmp_hop = c->tmp3.as<uint8_t>();
    if (k == 1) {
      LaunchScope ls(c, K_KHOP1_ROWS);
      const uint64_t blocks = std::min<uint64_t>((nq + K1_WARPS - 1) / K1_WARPS, (uint64_t)e->sm_count * 8);
      khop1_warp_kernel<<<(unsigned)std::max<uint64_t>(blocks, 1), K1_THREADS, 0, c->stream>>>(a, d_defer, d_ndefer);
    }
    {
      LaunchScope ls(c, use_large ? K_KHOP_LARGE : K_KHOP_GLOBAL);
      if (use_large)
        khop_large_kernel<<<(unsigned)grid, KL_THREADS, kl_smem, c->stream>>>(a, sc_queue, qcap, (uint32_t)(kl_smem / 5),
                                                                           d_defer, d_ndefer);
      else
        khop_global_kernel<<<(unsigned)grid, KG_THREADS, kg_smem, c->stream>>>(a, sc_seen, sc_rank, sc_queue, seen_words,
                                                                             qcap, d_defer, d_ndefer);
    }
    {
      LaunchScope ls(c, K_SCAN_COUNTS);
      scan_counts_kernel<<<(unsigned)std::min<uint64_t>(scan_tiles_all, (uint64_t)e->sm_count * 4), SC_THREADS, 0, c->stream>>>(
          a.counts, nq, a.out_off, &c->d_counters->aux[2], c->status.as<uint64_t>(), c->d_counters);
    }
    if (cap) {
      LaunchScope ls(c, K_KHOP_EMIT);
      // the row total is only known on the device: one resident wave, warps stride over the output
      khop_gather_kernel<<<(unsigned)e->sm_count * 8, 256, 0, c->stream>>>(nq, a.tmp_start, a.out_off, a.tmp_ids,
                                                                         a.tmp_hop, a.out_ids, a.out_hop, cap);
    }
  }
  DMG_CUDA(cudaGetLastError());
  DMG_TRY(fetch_counters(c));
  return DMG_OK;
}

static int32_t check_khop_args(dmg_graph* g, int mode, int k, uint64_t nq, const void* q_off, const void* out_off,
                               uint64_t* needed) {
  if (!g || !needed || (nq && !q_off) || !out_off) {
    set_error("dmg_neighbors_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (mode < DMG_OUT || mode > DMG_BOTH) {
    set_error("dmg_neighbors_batch: mode %d is not DMG_OUT/IN/BOTH", mode);
    return DMG_EINVAL;
  }
  if (k < 1 || k > DMG_MAX_HOPS) {
    set_error("dmg_neighbors_batch: k=%d outside 1..%d", k, DMG_MAX_HOPS);
    return DMG_EINVAL;
  }
  if (nq >= (1ull << 40)) {
    set_error("dmg_neighbors_batch: too many queries");
    return DMG_EINVAL;
  }
  return DMG_OK;
}

}  // namespace dmg

using namespace dmg;

extern "C" {

int32_t dmg_neighbors_batch_dev(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources, uint64_t nq,
                                const uint32_t* d_q_off, const uint32_t* d_q_src, uint64_t* d_out_off,
                                uint32_t* d_out_ids, uint8_t* d_out_hop, uint64_t cap, uint64_t* needed) {
  DMG_TRY(check_khop_args(g, mode, k, nq, d_q_off, d_out_off, needed));
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  *needed = 0;
  if (nq == 0) {
    DMG_CUDA(cudaMemsetAsync(d_out_off, 0, sizeof(uint64_t), c->stream));
    return c->finish();
  }
  DMG_TRY(run_khop(c, g, mode, k, exclude_sources, nq, d_q_off, d_q_src, d_out_off, d_out_ids, d_out_hop,
                   (d_out_ids && d_out_hop) ? cap : 0));
  c->stats.d2h_bytes += sizeof(DevCounters);
  DMG_TRY(c->finish());
  *needed = c->h_counters->total;
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = c->h_counters->rows;
  c->stats.results = c->h_counters->total;
  publish_stats(c->stats);
  if (*needed > cap) {
    set_error("dmg_neighbors_batch_dev: capacity %llu < needed %llu", (unsigned long long)cap,
              (unsigned long long)*needed);
    return DMG_ECAP;
  }
  return DMG_OK;
}

// ---- dmg_neighbors_batch_rows: dense result rows (one hop byte per project vertex) ---------------
int32_t dmg_neighbors_batch_rows(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources, uint64_t nq,
                                 const uint32_t* q_off, const uint32_t* q_src, uint32_t* out_row_stride,
                                 uint32_t* out_win, uint8_t* out_rows, uint64_t cap_bytes, uint64_t* needed_bytes) {
  if (!out_row_stride || !needed_bytes) {
    set_error("dmg_neighbors_batch_rows: NULL argument");
    return DMG_EINVAL;
  }
  uint64_t dummy_off = 0;
  DMG_TRY(check_khop_args(g, mode, k, nq, q_off, &dummy_off, needed_bytes));
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  if (maxp > SMALL_WINDOW_MAX) {
    set_error("dmg_neighbors_batch_rows: dense rows need projects of <= %u vertices (largest has %u)", SMALL_WINDOW_MAX, maxp);
    return DMG_EINVAL;
  }
  const uint32_t WC = (maxp + 127u) & ~127u;
  *out_row_stride = WC;
  *needed_bytes = nq * (uint64_t)WC;
  if (nq == 0) return DMG_OK;
  if (!out_win || !out_rows || cap_bytes < *needed_bytes) {
    set_error("dmg_neighbors_batch_rows: capacity %llu < needed %llu bytes", (unsigned long long)cap_bytes,
              (unsigned long long)*needed_bytes);
    return DMG_ECAP;
  }
  const uint64_t n_src = q_off[nq];
  if (n_src && !q_src) {
    set_error("dmg_neighbors_batch_rows: q_src is NULL");
    return DMG_EINVAL;
  }
  Engine* e = g->eng;
  CtxGuard cg(e);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_TRY(c->q_off.reserve((nq + 1) * 4));
  DMG_TRY(c->q_src.reserve(n_src * 4));
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, q_off, (nq + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (n_src) DMG_CUDA(cudaMemcpyAsync(c->q_src.p, q_src, n_src * 4, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += (nq + 1) * 4 + n_src * 4;
  const uint64_t stage_budget = std::max<uint64_t>(std::min<uint64_t>(e->scratch_budget, 2ull << 30), WC);
  const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(nq, stage_budget / WC));
  DMG_TRY(c->scratch.reserve(chunk * WC));
  DMG_TRY(c->tmp0.reserve(nq * 4));
  DMG_TRY(c->tmp1.reserve(nq * 8));
  DMG_TRY(reset_counters(c));
  KhopArgs a;
  memset(&a, 0, sizeof(a));
  a.out_ptr = g->d_out_ptr;
  a.out_idx = g->d_out_idx;
  a.in_ptr = g->d_in_ptr;
  a.in_idx = g->d_in_idx;
  a.und_ptr = g->d_und_ptr;
  a.und_idx = g->d_und_idx;
  a.part_off = g->d_part_off;
  a.V = g->info.n_vertices;
  a.n_parts = g->info.n_parts;
  a.mode = mode;
  a.k = k;
  a.exclude = exclude_sources ? 1 : 0;
  a.nq = nq;
  a.q_off = c->q_off.as<uint32_t>();
  a.q_src = c->q_src.as<uint32_t>();
  a.ctr = c->d_counters;
  a.win_cap = WC;
  a.stage = c->scratch.as<uint8_t>();
  a.counts = c->tmp0.as<uint32_t>();
  a.qwin = c->tmp1.as<uint32_t>();
  int occ = 0;
  size_t smem = 0;
  DMG_TRY(with_small_kernel(WC, [&](auto kern, uint32_t wc) -> int32_t {
    smem = small_kernel_smem(wc);
    if (smem > 48 * 1024)
      DMG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, KS_THREADS, smem));
    return DMG_OK;
  }));
  for (uint64_t q0 = 0; q0 < nq; q0 += chunk) {
    const uint64_t q1 = std::min<uint64_t>(nq, q0 + chunk);
    a.q_begin = q0;
    a.q_end = q1;
    DMG_CUDA(cudaMemsetAsync(&c->d_counters->ticket, 0, 8, c->stream));
    DMG_CUDA(cudaMemsetAsync(a.stage, 0, (q1 - q0) * WC, c->stream));   // bytes past a project's end read 0
    const uint64_t grid = std::min<uint64_t>(q1 - q0, (uint64_t)e->sm_count * std::max(occ, 1));
    {
      LaunchScope ls(c, K_KHOP_WINDOWS);
      launch_windows(a, c->stream);
    }
    {
      LaunchScope ls(c, K_KHOP_SMALL);
      with_small_kernel(WC, [&](auto kern, uint32_t) -> int32_t {
        kern<<<(unsigned)grid, KS_THREADS, smem, c->stream>>>(a);
        return DMG_OK;
      });
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(out_rows + q0 * WC, a.stage, (q1 - q0) * WC, cudaMemcpyDeviceToHost, c->stream));
    if (q1 < nq) DMG_CUDA(cudaStreamSynchronize(c->stream));   // the staging rows are reused by the next chunk
  }
  DMG_CUDA(cudaMemcpyAsync(out_win, a.qwin, nq * 8, cudaMemcpyDeviceToHost, c->stream));
  DMG_TRY(fetch_counters(c));
  c->stats.d2h_bytes += nq * (uint64_t)WC + nq * 8 + sizeof(DevCounters);
  DMG_TRY(c->finish());
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = c->h_counters->rows;
  publish_stats(c->stats);
  return DMG_OK;
}

// ---- dmg_neighbors_batch_bits: bit rows (binary-coded hop planes over the project's vertices) --------
int32_t dmg_neighbors_batch_bits(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources, uint64_t nq,
                                 const uint32_t* q_off, const uint32_t* q_src, uint32_t* out_planes,
                                 uint32_t* out_plane_words, uint32_t* out_win, uint32_t* out_counts, uint32_t* out_bits,
                                 uint64_t cap_bytes, uint64_t* needed_bytes) {
  if (!out_planes || !out_plane_words || !needed_bytes) {
    set_error("dmg_neighbors_batch_bits: NULL argument");
    return DMG_EINVAL;
  }
  uint64_t dummy_off = 0;
  DMG_TRY(check_khop_args(g, mode, k, nq, q_off, &dummy_off, needed_bytes));
  const int w64 = bit_parallel_w64(g, k);
  if (!w64) {
    set_error("dmg_neighbors_batch_bits: bit rows need projects of <= %u vertices and k <= 7 (largest project %u, k = %d)",
              MB_MAX_WINDOW, g->info.max_part_vertices, k);
    return DMG_EINVAL;
  }
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  const uint32_t WB = ((maxp + 127u) & ~127u) / 32;
  const int np = hop_planes(k);
  const uint64_t row_bytes = (uint64_t)np * WB * 4;
  *out_planes = (uint32_t)np;
  *out_plane_words = WB;
  *needed_bytes = nq * row_bytes;
  if (nq == 0) return DMG_OK;
  if (!out_win || !out_counts || !out_bits || cap_bytes < *needed_bytes) {
    set_error("dmg_neighbors_batch_bits: capacity %llu < needed %llu bytes", (unsigned long long)cap_bytes,
              (unsigned long long)*needed_bytes);
    return DMG_ECAP;
  }
  const uint64_t n_src = q_off[nq];
  if (n_src && !q_src) {
    set_error("dmg_neighbors_batch_bits: q_src is NULL");
    return DMG_EINVAL;
  }
  Engine* e = g->eng;
  CtxGuard cg(e);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_TRY(c->q_off.reserve((nq + 1) * 4));
  DMG_TRY(c->q_src.reserve(n_src * 4));
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, q_off, (nq + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (n_src) DMG_CUDA(cudaMemcpyAsync(c->q_src.p, q_src, n_src * 4, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += (nq + 1) * 4 + n_src * 4;
  const uint64_t budget = std::max<uint64_t>(std::min<uint64_t>(e->scratch_budget, 2ull << 30), row_bytes);
  const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(nq, budget / row_bytes));
  const MbLayout L(chunk, g->info.n_parts, 64u * (uint32_t)w64);
  DMG_TRY(c->scratch.reserve(chunk * row_bytes));
  DMG_TRY(c->tmp0.reserve(nq * 4));
  DMG_TRY(c->tmp1.reserve(nq * 8));
  DMG_TRY(c->tmp2.reserve(L.bytes));
  DMG_TRY(reset_counters(c));
  for (uint64_t q0 = 0; q0 < nq; q0 += chunk) {
    const uint64_t q1 = std::min<uint64_t>(nq, q0 + chunk);
    if (q0) DMG_CUDA(cudaMemsetAsync(&c->d_counters->ticket, 0, 8, c->stream));
    DMG_TRY(launch_bit_traversal(c, g, w64, mode, k, exclude_sources, q0, q1, c->q_off.as<uint32_t>(),
                                 c->q_src.as<uint32_t>(), c->scratch.as<uint32_t>(), c->tmp0.as<uint32_t>(),
                                 c->tmp1.as<uint32_t>(), c->tmp2.as<unsigned char>(), c->d_counters, c->stream));
    {
      LaunchScope ls(c, K_MISC);
      mb_zero_orphan_rows_kernel<<<(unsigned)((q1 - q0 + 7) / 8), 256, 0, c->stream>>>(
          c->tmp2.as<uint32_t>(), q1 - q0, c->scratch.as<uint32_t>(), (uint32_t)(row_bytes / 4));
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(reinterpret_cast<unsigned char*>(out_bits) + q0 * row_bytes, c->scratch.p, (q1 - q0) * row_bytes,
                             cudaMemcpyDeviceToHost, c->stream));
    if (q1 < nq) DMG_CUDA(cudaStreamSynchronize(c->stream));   // the rows are reused by the next chunk
  }
  DMG_CUDA(cudaMemcpyAsync(out_win, c->tmp1.p, nq * 8, cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaMemcpyAsync(out_counts, c->tmp0.p, nq * 4, cudaMemcpyDeviceToHost, c->stream));
  DMG_TRY(fetch_counters(c));
  c->stats.d2h_bytes += nq * row_bytes + nq * 12 + sizeof(DevCounters);
  DMG_TRY(c->finish());
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = c->h_counters->rows;
  publish_stats(c->stats);
  return DMG_OK;
}

// ---- dmg_neighbors_batch_dev_static: the k-hop batch without a single allocation ---------------
// Scratch for the small-project kernel chain (counters, scan status words, row sizes, per-query
// window, staging rows) lives in this file-scope device array instead of cudaMalloc'ed buffers.
// Every GPU that loads the library carries the 16 MiB until the process exits.
static constexpr size_t kStaticScratchBytes = 16u << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
// In host code `g_static_scratch` is only a handle for the runtime, not a device address; the
// real address is looked up once per device (with that device current) and cached here.
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[64] = {nullptr};

static int32_t static_scratch_for_device(int dev, unsigned char** out) {
  if (dev < 0 || dev >= 64) {
    set_error("dmg_neighbors_batch_dev_static: device ordinal %d not supported", dev);
    return DMG_EINVAL;
  }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* p = nullptr;
    DMG_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    g_static_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[dev];
  return DMG_OK;
}

static inline size_t pad256(size_t x) { return (x + 255) & ~size_t(255); }

// The part of the static entry point that runs with the engine's device current and may have queued
// work on the default stream when it fails; the wrapper below waits for that work and puts the
// caller's device back whatever this returns.
static int32_t static_khop_on_device(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources, uint64_t nq,
                                     const uint32_t* d_q_off, const uint32_t* d_q_src, uint64_t* d_out_off,
                                     uint32_t* d_out_ids, uint8_t* d_out_hop, uint64_t cap, uint64_t* needed) {
  Engine* e = g->eng;
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  if (nq == 0) {
    DMG_CUDA(cudaMemsetAsync(d_out_off, 0, sizeof(uint64_t), 0));
    DMG_CUDA(cudaStreamSynchronize(0));
    return DMG_OK;
  }
  const uint32_t WC = (maxp + 127u) & ~127u;
  const uint64_t tiles = (nq + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
  // layout: [counters][scan status][row sizes][(lo, wn) per query][staging rows]
  const size_t o_status = pad256(sizeof(DevCounters));
  const size_t o_counts = o_status + pad256(tiles * sizeof(uint64_t));
  const size_t o_qwin = o_counts + pad256(nq * 4);
  const size_t o_stage = o_qwin + pad256(nq * 8);
  const unsigned long long want = (unsigned long long)o_stage + (unsigned long long)nq * WC;
  if (nq > (1ull << 32) || want > kStaticScratchBytes) {
    set_error("dmg_neighbors_batch_dev_static: %llu queries over %u-vertex projects need %llu bytes of scratch; "
              "the static scratch holds %llu (split the batch or use dmg_neighbors_batch_dev)",
              (unsigned long long)nq, WC, want, (unsigned long long)kStaticScratchBytes);
    return DMG_EINVAL;
  }
  unsigned char* base = nullptr;
  DMG_TRY(static_scratch_for_device(e->device, &base));
  DevCounters* ctr = reinterpret_cast<DevCounters*>(base);
  // the array keeps the previous call's contents: counters and status words must start at zero
  DMG_CUDA(cudaMemsetAsync(base, 0, o_counts, 0));
  KhopArgs a;
  a.out_ptr = g->d_out_ptr;
  a.out_idx = g->d_out_idx;
  a.in_ptr = g->d_in_ptr;
  a.in_idx = g->d_in_idx;
  a.und_ptr = g->d_und_ptr;
  a.und_idx = g->d_und_idx;
  a.part_off = g->d_part_off;
  a.V = g->info.n_vertices;
  a.n_parts = g->info.n_parts;
  a.mode = mode;
  a.k = k;
  a.exclude = exclude_sources ? 1 : 0;
  a.nq = nq;
  a.q_off = d_q_off;
  a.q_src = d_q_src;
  a.out_off = reinterpret_cast<unsigned long long*>(d_out_off);
  a.out_ids = d_out_ids;
  a.out_hop = d_out_hop;
  a.cap = (d_out_ids && d_out_hop) ? cap : 0;
  a.status = reinterpret_cast<uint64_t*>(base + o_status);
  a.ctr = ctr;
  a.win_cap = WC;
  a.q_begin = 0;
  a.q_end = nq;
  a.stage = base + o_stage;
  a.counts = reinterpret_cast<uint32_t*>(base + o_counts);
  a.qwin = reinterpret_cast<uint32_t*>(base + o_qwin);
  a.tmp_start = nullptr;
  a.tmp_ids = nullptr;
  a.tmp_hop = nullptr;
  int occ = 0;
  size_t smem = 0;
  DMG_TRY(with_small_kernel(WC, [&](auto kern, uint32_t wc) -> int32_t {
    smem = small_kernel_smem(wc);
    if (smem > 48 * 1024)
      DMG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, KS_THREADS, smem));
    return DMG_OK;
  }));
  const uint64_t grid = std::min<uint64_t>(nq, (uint64_t)e->sm_count * std::max(occ, 1));
  launch_windows(a, 0);
  with_small_kernel(WC, [&](auto kern, uint32_t) -> int32_t {
    kern<<<(unsigned)grid, KS_THREADS, smem, 0>>>(a);
    return DMG_OK;
  });
  scan_counts_kernel<<<(unsigned)std::min<uint64_t>(tiles, (uint64_t)e->sm_count * 4), SC_THREADS, 0, 0>>>(
      a.counts, nq, a.out_off, &ctr->aux[2], a.status, ctr);
  khop_emit_kernel<<<(unsigned)std::min<uint64_t>((nq + 7) / 8, (uint64_t)e->sm_count * EM_GRID_PER_SM), EM_THREADS, 0, 0>>>(
      a.stage, WC, 0, nq, a.qwin, a.out_off, a.out_ids, a.out_hop, a.cap);
  DMG_CUDA(cudaGetLastError());
  unsigned long long total = 0;
  DMG_CUDA(cudaMemcpy(&total, &ctr->total, sizeof(total), cudaMemcpyDeviceToHost));   // waits for the kernels
  *needed = total;
  if (total > cap) {
    set_error("dmg_neighbors_batch_dev_static: capacity %llu < needed %llu", (unsigned long long)cap, total);
    return DMG_ECAP;
  }
  return DMG_OK;
}

int32_t dmg_neighbors_batch_dev_static(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources, uint64_t nq,
                                       const uint32_t* d_q_off, const uint32_t* d_q_src, uint64_t* d_out_off,
                                       uint32_t* d_out_ids, uint8_t* d_out_hop, uint64_t cap, uint64_t* needed) {
  DMG_TRY(check_khop_args(g, mode, k, nq, d_q_off, d_out_off, needed));
  *needed = 0;
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  if (maxp > SMALL_WINDOW_MAX) {
    set_error("dmg_neighbors_batch_dev_static: largest project has %u vertices; the static-scratch variant "
              "covers projects up to %u (use dmg_neighbors_batch_dev)", maxp, SMALL_WINDOW_MAX);
    return DMG_EINVAL;
  }
  // The work runs on the engine's device; the caller's current device is put back before returning.
  int prev_dev = -1;
  cudaGetDevice(&prev_dev);
  DMG_CUDA(cudaSetDevice(g->eng->device));
  const int32_t rc = static_khop_on_device(g, mode, k, exclude_sources, nq, d_q_off, d_q_src, d_out_off, d_out_ids,
                                           d_out_hop, cap, needed);
  if (rc != DMG_OK && rc != DMG_ECAP) {
    // a failure after the first launch leaves kernels running over the shared array: wait for them,
    // so that the caller may retry or issue another call at once
    cudaStreamSynchronize(0);
    cudaGetLastError();
  }
  if (prev_dev >= 0 && prev_dev != g->eng->device) cudaSetDevice(prev_dev);
  return rc;
}

// dmg_neighbors_batch on small projects, large batches: the packed (id, hop) rows are ~20x the bytes of the
// bit rows they come from (config 2: 652 MB vs 34 MB per batch), and PCIe is what the host-pointer call waits
// for.  So the batch crosses the bus as bit rows and is expanded by the host's threads (bits_decode.cc)
// straight into the caller's arrays (very large batches in chunks, the expansion of one chunk running while
// the next is traversed and copied).  Same rows, offsets and capacity answer as the device-side emit.
static int32_t neighbors_batch_via_bits(CallCtx* c, dmg_graph* g, int w64, int mode, int k, int exclude, uint64_t nq,
                                        const uint32_t* q_off, const uint32_t* q_src, uint64_t n_src, uint64_t* out_off,
                                        uint32_t* out_ids, uint8_t* out_hop, uint64_t cap, uint64_t* needed) {
  const uint32_t maxp = std::max<uint32_t>(g->info.max_part_vertices, 1);
  const uint32_t WB = ((maxp + 127u) & ~127u) / 32;
  const int np = hop_planes(k);
  const size_t row_words = (size_t)np * WB, row_bytes = row_words * 4;
  // one chunk per 262,144 queries: with smaller chunks the expansion of one overlapped the copy of the next,
  // but eight thread start-ups and eight short expansions cost more than the 1 ms of copy they hid
  // (config 2: 6.1 ms in eight chunks, 5.1 ms in one)
  const uint64_t n_chunks = std::max<uint64_t>(1, std::min<uint64_t>(8, nq / 262144));
  const uint64_t chunk = (nq + n_chunks - 1) / n_chunks;
  const MbLayout L(chunk, g->info.n_parts, 64u * (uint32_t)w64);
  DMG_TRY(c->q_off.reserve((nq + 1) * 4));
  DMG_TRY(c->q_src.reserve(n_src * 4));
  DMG_TRY(c->scratch.reserve(nq * row_bytes));
  DMG_TRY(c->tmp0.reserve(nq * 4));
  DMG_TRY(c->tmp1.reserve(nq * 8));
  DMG_TRY(c->tmp2.reserve(L.bytes));
  const size_t o_win = (nq * row_bytes + 255) & ~size_t(255), o_cnt = (o_win + nq * 8 + 255) & ~size_t(255);
  DMG_TRY(c->reserve_pinned(o_cnt + nq * 4));
  unsigned char* pin = static_cast<unsigned char*>(c->h_pin);
  uint32_t* h_bits = reinterpret_cast<uint32_t*>(pin);
  uint32_t* h_win = reinterpret_cast<uint32_t*>(pin + o_win);
  uint32_t* h_cnt = reinterpret_cast<uint32_t*>(pin + o_cnt);
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, q_off, (nq + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (n_src) DMG_CUDA(cudaMemcpyAsync(c->q_src.p, q_src, n_src * 4, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += (nq + 1) * 4 + n_src * 4;
  DMG_TRY(reset_counters(c));
  std::vector<cudaEvent_t> done;
  for (uint64_t q0 = 0; q0 < nq; q0 += chunk) {
    const uint64_t q1 = std::min<uint64_t>(nq, q0 + chunk);
    if (q0) DMG_CUDA(cudaMemsetAsync(&c->d_counters->ticket, 0, 8, c->stream));
    DMG_TRY(launch_bit_traversal(c, g, w64, mode, k, exclude, q0, q1, c->q_off.as<uint32_t>(), c->q_src.as<uint32_t>(),
                                 c->scratch.as<uint32_t>() + q0 * row_words, c->tmp0.as<uint32_t>(),
                                 c->tmp1.as<uint32_t>(), c->tmp2.as<unsigned char>(), c->d_counters, c->stream));
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(h_cnt + q0, c->tmp0.as<uint32_t>() + q0, (q1 - q0) * 4, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaMemcpyAsync(h_win + 2 * q0, c->tmp1.as<uint32_t>() + 2 * q0, (q1 - q0) * 8, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaMemcpyAsync(h_bits + q0 * row_words, c->scratch.as<uint32_t>() + q0 * row_words, (q1 - q0) * row_bytes,
                             cudaMemcpyDeviceToHost, c->stream));
    cudaEvent_t ev = c->get_event();
    DMG_CUDA(cudaEventRecord(ev, c->stream));
    done.push_back(ev);
  }
  DMG_TRY(fetch_counters(c));
  c->stats.d2h_bytes += nq * row_bytes + nq * 12 + sizeof(DevCounters);
  // the host follows the device chunk by chunk (a row of a query without a valid source has count 0 and is
  // never read, so those rows need no zeroing here)
  uint64_t total = 0;
  bool fits = out_ids && out_hop;
  size_t ci = 0;
  int32_t rc = DMG_OK;
  for (uint64_t q0 = 0; q0 < nq; q0 += chunk, ++ci) {
    const uint64_t q1 = std::min<uint64_t>(nq, q0 + chunk);
    if (cudaEventSynchronize(done[ci]) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaEventSynchronize", __FILE__, __LINE__);
    if (rc != DMG_OK) break;
    for (uint64_t q = q0; q < q1; ++q) {
      out_off[q] = total;
      total += h_cnt[q];
    }
    out_off[q1] = total;
    if (fits && total <= cap) bits_decode_rows((uint32_t)np, WB, q0, q1, h_win, h_cnt, h_bits, out_off, out_ids, out_hop, 0);
    else fits = false;
  }
  for (cudaEvent_t ev : done) c->ev_pool.push_back(ev);
  DMG_TRY(rc);
  *needed = total;
  c->stats.results = total;
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = c->h_counters->rows;
  DMG_TRY(c->finish());
  if (total > cap || (total && !fits)) {
    set_error("dmg_neighbors_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    return DMG_ECAP;
  }
  return DMG_OK;
}

int32_t dmg_neighbors_batch(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources, uint64_t nq,
                            const uint32_t* q_off, const uint32_t* q_src, uint64_t* out_off, uint32_t* out_ids,
                            uint8_t* out_hop, uint64_t cap, uint64_t* needed) {
  DMG_TRY(check_khop_args(g, mode, k, nq, q_off, out_off, needed));
  *needed = 0;
  out_off[0] = 0;
  if (nq == 0) return DMG_OK;
  const uint64_t n_src = q_off[nq];
  if (n_src && !q_src) {
    set_error("dmg_neighbors_batch: q_src is NULL");
    return DMG_EINVAL;
  }
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  if (const int w64 = bit_parallel_w64(g, k)) {
    // fetch calls of large batches on small projects cross the bus as bit rows (DMG_HOST_PACKED=1: the packed
    // rows of the device-side emit instead, for A/B runs)
    const uint32_t WBq = ((std::max<uint32_t>(g->info.max_part_vertices, 1) + 127u) & ~127u) / 32;
    const uint64_t bits_bytes = nq * (uint64_t)hop_planes(k) * WBq * 4;
    const bool host_packed = getenv("DMG_HOST_PACKED") != nullptr;
    if (!host_packed && out_ids && out_hop && cap && nq >= 2048 && bits_bytes <= (1ull << 30) &&
        bits_bytes <= g->eng->scratch_budget)
      return neighbors_batch_via_bits(c, g, w64, mode, k, exclude_sources, nq, q_off, q_src, n_src, out_off, out_ids,
                                      out_hop, cap, needed);
  }
  // a row can never exceed its project, so the device buffers are bounded even for a huge cap
  const uint64_t row_max = g->info.max_part_vertices;
  uint64_t dev_cap = (!out_ids || !out_hop) ? 0 : cap;
  if (row_max && nq <= UINT64_MAX / std::max<uint64_t>(row_max, 1)) dev_cap = std::min<uint64_t>(dev_cap, nq * row_max);
  DMG_TRY(c->q_off.reserve((nq + 1) * 4));
  DMG_TRY(c->q_src.reserve(n_src * 4));
  DMG_TRY(c->out_off.reserve((nq + 1) * 8));
  DMG_TRY(c->out_ids.reserve(dev_cap * 4));
  DMG_TRY(c->out_hop.reserve(dev_cap));
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, q_off, (nq + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (n_src) DMG_CUDA(cudaMemcpyAsync(c->q_src.p, q_src, n_src * 4, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += (nq + 1) * 4 + n_src * 4;
  DMG_TRY(run_khop(c, g, mode, k, exclude_sources, nq, c->q_off.as<uint32_t>(), c->q_src.as<uint32_t>(),
                   c->out_off.as<uint64_t>(), c->out_ids.as<uint32_t>(), c->out_hop.as<uint8_t>(), dev_cap));
  DMG_CUDA(cudaMemcpyAsync(out_off, c->out_off.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  const uint64_t total = c->h_counters->total;
  c->stats.d2h_bytes += (nq + 1) * 8 + sizeof(DevCounters);
  *needed = total;
  int32_t rc = DMG_OK;
  if (total > cap || (total && (!out_ids || !out_hop))) {
    set_error("dmg_neighbors_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    rc = DMG_ECAP;
  } else if (total) {
    DMG_CUDA(cudaMemcpyAsync(out_ids, c->out_ids.p, total * 4, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaMemcpyAsync(out_hop, c->out_hop.p, total, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += total * 5;
  }
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = c->h_counters->rows;
  c->stats.results = total;
  DMG_TRY(c->finish());
  return rc;
}

int32_t dmg_adjacency_batch(dmg_graph* g, int32_t dir, uint64_t n, const uint32_t* ids, uint64_t* out_off,
                            uint32_t* out_ids, uint64_t cap, uint64_t* needed) {
  if (!g || !out_off || !needed || (n && !ids)) {
    set_error("dmg_adjacency_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (dir != DMG_OUT && dir != DMG_IN) {
    set_error("dmg_adjacency_batch: dir %d is not DMG_OUT/DMG_IN", dir);
    return DMG_EINVAL;
  }
  *needed = 0;
  out_off[0] = 0;
  if (n == 0) return DMG_OK;
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  const uint64_t n_tiles = (n + ADJ_THREADS - 1) / ADJ_THREADS;
  DMG_TRY(c->q_src.reserve(n * 4));
  DMG_TRY(c->out_off.reserve((n + 1) * 8));
  DMG_TRY(c->status.reserve(n_tiles * 8));
  DMG_TRY(c->tmp0.reserve(n * 16));   // (start, len, out offset) of rows left to adjacency_long_kernel
  DMG_TRY(c->tmp1.reserve(n * 4 + 16));   // their first chunk
  DMG_CUDA(cudaMemcpyAsync(c->q_src.p, ids, n * 4, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += n * 4;
  const uint32_t* ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
  const uint32_t* idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
  const uint64_t grid = std::min<uint64_t>(n_tiles, (uint64_t)g->eng->sm_count * 8);
  // One pass when the caller's capacity is a sane device allocation; otherwise a sizing pass
  // (capacity 0: offsets only) followed by the gather into an exactly sized buffer.
  const uint64_t cap_eff = out_ids ? cap : 0;
  const bool one_pass = cap_eff * 4 <= std::max<uint64_t>(256ull << 20, c->out_ids.cap);
  uint64_t total = 0;
  for (int pass = one_pass ? 1 : 0; pass < 2; ++pass) {
    const uint64_t use_cap = pass == 0 ? 0 : (one_pass ? cap_eff : std::min<uint64_t>(cap_eff, total));
    if (use_cap) DMG_TRY(c->out_ids.reserve(use_cap * 4));
    DMG_CUDA(cudaMemsetAsync(c->status.p, 0, n_tiles * 8, c->stream));
    DMG_TRY(reset_counters(c));
    {
      LaunchScope ls(c, K_ADJ_GATHER);
      adjacency_kernel<<<(unsigned)grid, ADJ_THREADS, 0, c->stream>>>(
          ptr, idx, g->info.n_vertices, n, c->q_src.as<uint32_t>(), c->out_off.as<unsigned long long>(),
          c->out_ids.as<uint32_t>(), use_cap, c->status.as<uint64_t>(), c->d_counters,
          c->tmp0.as<unsigned long long>());
    }
    if (use_cap) {
      LaunchScope ls(c, K_ADJ_GATHER, true, 2);
      adjacency_long_scan_kernel<<<1, 1024, 0, c->stream>>>(c->tmp0.as<unsigned long long>(), c->d_counters,
                                                            c->tmp1.as<uint32_t>());
      adjacency_long_kernel<<<g->eng->sm_count * 8, 256, 0, c->stream>>>(idx, c->tmp0.as<unsigned long long>(),
                                                                       c->tmp1.as<uint32_t>(), c->d_counters,
                                                                       c->out_ids.as<uint32_t>(), use_cap);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_TRY(fetch_counters(c));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    total = c->h_counters->total;
    if (pass == 0 && (total > cap_eff || total == 0)) break;
  }
  DMG_CUDA(cudaMemcpyAsync(out_off, c->out_off.p, (n + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  c->stats.d2h_bytes += (n + 1) * 8 + sizeof(DevCounters);
  *needed = total;
  int32_t rc = DMG_OK;
  if (total > cap || (total && !out_ids)) {
    set_error("dmg_adjacency_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    rc = DMG_ECAP;
  } else if (total) {
    DMG_CUDA(cudaMemcpyAsync(out_ids, c->out_ids.p, total * 4, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += total * 4;
  }
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = n;
  c->stats.results = total;
  DMG_TRY(c->finish());
  return rc;
}

int32_t dmg_entry_points(dmg_graph* g, uint32_t* out_ids, uint32_t cap, uint32_t* needed) {
  if (!g || !needed) {
    set_error("dmg_entry_points: NULL argument");
    return DMG_EINVAL;
  }
  // the ordered entry list is tiny, immutable and kept on both sides; served from the handle
  *needed = (uint32_t)g->h_entry.size();
  if (*needed > cap || (*needed && !out_ids)) return DMG_ECAP;
  if (*needed) memcpy(out_ids, g->h_entry.data(), *needed * sizeof(uint32_t));
  return DMG_OK;
}

int32_t dmg_is_entry_point_batch(dmg_graph* g, uint64_t n, const uint32_t* ids, uint8_t* out_flags) {
  if (!g || (n && (!ids || !out_flags))) {
    set_error("dmg_is_entry_point_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (!n) return DMG_OK;
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_TRY(c->q_src.reserve(n * 4));
  DMG_TRY(c->out_hop.reserve(n));
  DMG_CUDA(cudaMemcpyAsync(c->q_src.p, ids, n * 4, cudaMemcpyHostToDevice, c->stream));
  {
    LaunchScope ls(c, K_ENTRY_FLAGS);
    entry_flags_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(g->d_entry_bits, g->info.n_vertices, n,
                                                                          c->q_src.as<uint32_t>(),
                                                                          c->out_hop.as<uint8_t>());
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(out_flags, c->out_hop.p, n, cudaMemcpyDeviceToHost, c->stream));
  c->stats.h2d_bytes += n * 4;
  c->stats.d2h_bytes += n;
  return c->finish();
}

}  // extern "C"
