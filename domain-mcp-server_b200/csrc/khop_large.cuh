// k-hop on projects too large for the shared-memory sets of khop_small / khop_msbfs but whose CLAIM
// BITMAP still fits one SM's shared memory (up to ~1.4 M vertices: BASELINE config 4's 1 M-vertex
// graph).  Included by traverse.cu.
//
//   khop1_warp_kernel   k = 1: a result row is the union of the sources' adjacency rows -- no traversal
//                       state at all.  One warp per query: rows are copied into a per-warp shared-memory
//                       buffer, ordered by a bitonic network, de-duplicated in place and appended to the
//                       cursor-ordered temporaries.  Queries whose candidate rows do not fit the buffer
//                       (hub rows) or with more than 32 sources are put on a list for the CTA kernel.
//   khop_large_kernel   k >= 2 (and the deferred k = 1 queries): one 1024-thread CTA per SM and query.
//                       khop_global_kernel keeps its claim bitmap in global memory, which makes every
//                       adjacency entry a dependent L2 round trip (load id -> load word -> atomicOr);
//                       here the bitmap lives in shared memory, the id loads of four entries are in
//                       flight per lane, and the ascending-id order of the row comes from ranking the
//                       discovered vertices inside the bitmap (per-quad popcount prefix, also shared
//                       memory) instead of sorting them.
// Both leave (tmp_start[q], counts[q]) and rows in cursor order; scan_counts_kernel + khop_gather_kernel
// put them into query order exactly as for khop_global_kernel.
#pragma once

namespace dmg {

constexpr int K1_THREADS = 256, K1_WARPS = K1_THREADS / 32;
constexpr uint32_t K1_CAP = 1024;     // candidate entries one warp orders in shared memory

__global__ void __launch_bounds__(K1_THREADS) khop1_warp_kernel(const KhopArgs a, uint32_t* defer, uint32_t* n_defer) {
  __shared__ uint32_t s_buf[K1_WARPS][K1_CAP];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  uint32_t* buf = s_buf[wib];
  const uint32_t* __restrict__ fptr = a.mode == DMG_BOTH ? a.und_ptr : (a.mode == DMG_OUT ? a.out_ptr : a.in_ptr);
  const uint32_t* __restrict__ fidx = a.mode == DMG_BOTH ? a.und_idx : (a.mode == DMG_OUT ? a.out_idx : a.in_idx);
  const unsigned long long warp0 = (blockIdx.x * (unsigned long long)K1_THREADS + threadIdx.x) >> 5;
  const unsigned long long n_warps = ((unsigned long long)gridDim.x * K1_THREADS) >> 5;
  unsigned long long te = 0, rows = 0;
  for (unsigned long long q = warp0; q < a.nq; q += n_warps) {
    const uint32_t qb = a.q_off[q], qe = a.q_off[q + 1];
    bool deferred = qe - qb > 32;
    uint32_t lo = 0, hi = 0;
    if (!deferred) {
      if (lane == 0) find_window(a, qb, qe, &lo, &hi);
      lo = __shfl_sync(FULL, lo, 0);
      hi = __shfl_sync(FULL, hi, 0);
    }
    // lane i holds source i if it is the first occurrence of a vertex of the window
    uint32_t s = 0xFFFFFFFFu, b = 0, n = 0;
    if (!deferred && qb + lane < qe) {
      const uint32_t x = a.q_src[qb + lane];
      if (x >= lo && x < hi) s = x;
    }
    if (!deferred) {
      const unsigned same = __match_any_sync(FULL, s);
      if (s != 0xFFFFFFFFu && (same & lanemask_lt())) s = 0xFFFFFFFFu;
      if (s != 0xFFFFFFFFu) {
        b = fptr[s];
        n = fptr[s + 1] - b;
      }
    }
    const uint32_t inc = warp_incl_scan_u32(n);
    const uint32_t T = __shfl_sync(FULL, inc, 31);
    const unsigned src_mask = __ballot_sync(FULL, s != 0xFFFFFFFFu);
    if (deferred || T > K1_CAP) {
      if (lane == 0) defer[atomicAdd(n_defer, 1u)] = (uint32_t)q;
      continue;
    }
    te += n;
    rows += s != 0xFFFFFFFFu;
    if (T == 0) {   // no source of the window, or sources without rows
      if (lane == 0) {
        a.tmp_start[q] = 0;
        a.counts[q] = 0;
      }
      continue;
    }
    uint32_t P = 32;
    while (P < T) P <<= 1;
    // rows -> buffer, row after row (coalesced); the tail up to the next power of two sorts last
    for (unsigned m = src_mask; m; m &= m - 1) {
      const int j = __ffs(m) - 1;
      const uint32_t bj = __shfl_sync(FULL, b, j), nj = __shfl_sync(FULL, n, j), oj = __shfl_sync(FULL, inc - n, j);
      for (uint32_t t = lane; t < nj; t += 32) buf[oj + t] = fidx[bj + t];
    }
    for (uint32_t t = T + lane; t < P; t += 32) buf[t] = 0xFFFFFFFFu;
    __syncwarp();
    for (uint32_t size = 2; size <= P; size <<= 1)
      for (uint32_t stride = size >> 1; stride; stride >>= 1) {
        for (uint32_t t = lane; t < (P >> 1); t += 32) {
          const uint32_t i = ((t & ~(stride - 1)) << 1) | (t & (stride - 1)), j = i | stride;
          const uint32_t x = buf[i], y = buf[j];
          const bool up = (i & size) == 0;
          if ((x > y) == up) {
            buf[i] = y;
            buf[j] = x;
          }
        }
        __syncwarp();
      }
    // keep one copy of every id; excluded sources are not results (a source that another source points
    // to IS one when sources are kept: hop 1).  Compaction in place, 32 entries per step, in order.
    uint32_t total = 0;
    for (uint32_t t0 = 0; t0 < T; t0 += 32) {
      const uint32_t t = t0 + lane;
      uint32_t v = 0xFFFFFFFFu;
      bool keep = false;
      if (t < T) {
        v = buf[t];
        keep = t == 0 || buf[t - 1] != v;
      }
      if (a.exclude)
        for (unsigned m = src_mask; m; m &= m - 1) {
          const uint32_t sj = __shfl_sync(FULL, s, __ffs(m) - 1);   // every lane takes part, kept or not
          keep = keep && v != sj;
        }
      const unsigned km = __ballot_sync(FULL, keep);
      __syncwarp();
      if (keep) buf[total + __popc(km & lanemask_lt())] = v;
      total += __popc(km);
      __syncwarp();
    }
    unsigned long long at = 0;
    if (lane == 0) {
      at = atomicAdd(&a.ctr->aux[3], (unsigned long long)total);
      a.tmp_start[q] = at;
      a.counts[q] = total;
    }
    at = __shfl_sync(FULL, at, 0);
    for (uint32_t t = lane; t < total; t += 32)
      if (at + t < a.cap) {
        a.tmp_ids[at + t] = buf[t];
        a.tmp_hop[at + t] = 1;
      }
    __syncwarp();
  }
  te = warp_sum_u64(te);
  rows = warp_sum_u64(rows);
  if (lane == 0 && rows) {
    atomicAdd(&a.ctr->edges, te);
    atomicAdd(&a.ctr->aux[0], te);
    atomicAdd(&a.ctr->rows, rows);
  }
}

// ---------------------------------------------------------------------------------------------------
constexpr int KL_THREADS = 1024, KL_NW = KL_THREADS / 32;
constexpr int KL_ILP = 4;             // adjacency / queue entries in flight per lane (8 measured the same)
constexpr int KL_SEGR = 3;            // level boundaries kept in registers by the emit (k <= 4 needs no table walk)
// dynamic shared memory: claim bitmap (W words, W a multiple of 128) + one rank word per quad of bitmap words
static inline size_t large_kernel_smem(uint32_t maxp) {
  const size_t W = (((size_t)maxp + 31) / 32 + 127) & ~size_t(127);
  return W * 4 + (W / 4) * 4;
}

__global__ void __launch_bounds__(KL_THREADS, 1) khop_large_kernel(const KhopArgs a, uint32_t* sc_queue, size_t qcap,
                                                                    uint32_t W, const uint32_t* qlist,
                                                                    const uint32_t* qlist_n) {
  constexpr int THREADS = KL_THREADS, NW = KL_NW;
  extern __shared__ __align__(16) unsigned char kl_smem[];
  uint32_t* seen = reinterpret_cast<uint32_t*>(kl_smem);   // [W]
  uint32_t* rankq = seen + W;                               // [W / 4] rank of the quad's first bit, relative to its warp segment
  __shared__ unsigned long long s_q, s_base;
  __shared__ uint32_t s_lo, s_hi, s_tail;
  __shared__ uint32_t s_seg[DMG_MAX_HOPS + 2];
  __shared__ uint32_t s_wsum[NW + 1];
  __shared__ uint32_t s_segbase[NW];
  __shared__ uint2 s_row[THREADS];
  uint32_t* queue = sc_queue + (size_t)blockIdx.x * qcap;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t* __restrict__ fptr = a.mode == DMG_BOTH ? a.und_ptr : (a.mode == DMG_OUT ? a.out_ptr : a.in_ptr);
  const uint32_t* __restrict__ fidx = a.mode == DMG_BOTH ? a.und_idx : (a.mode == DMG_OUT ? a.out_idx : a.in_idx);
  const unsigned long long n_run = qlist ? (unsigned long long)*qlist_n : a.nq;
  unsigned long long te = 0, rows = 0;
  for (uint32_t i = tid; i < W; i += THREADS) seen[i] = 0;
  // quads of the bitmap are ranked per warp segment of QS quads (a power of two, so a quad finds its
  // segment with a shift)
  const uint32_t nquads_all = W >> 2;
  uint32_t qs_log = 0;
  while ((nquads_all + NW - 1) / NW > (1u << qs_log)) ++qs_log;
  __syncthreads();

  while (true) {
    if (tid == 0) s_q = atomicAdd(&a.ctr->ticket, 1ull);
    __syncthreads();
    const unsigned long long t = s_q;
    if (t >= n_run) break;
    const unsigned long long q = qlist ? qlist[t] : t;
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
    if (!a.exclude) {   // kept sources may be found again: only excluded sources stay claimed (CCS:1563-1578 iterated)
      for (uint32_t i = tid; i < ns; i += THREADS) {
        const uint32_t v = queue[i];
        atomicAnd(&seen[v >> 5], ~(1u << (v & 31)));
      }
      __syncthreads();
    }
    if (tid == 0) s_seg[0] = ns;
    int nlev = 0;
    for (int lvl = 1; lvl <= a.k && head < tail; ++lvl) {
      for (uint32_t c0 = head; c0 < tail; c0 += THREADS) {
        const uint32_t i = c0 + tid;
        uint32_t b = 0, n = 0;
        if (i < tail) {
          const uint32_t u = lo + __ldcg(&queue[i]);
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
        // every warp takes a contiguous share of the T flattened entries, KL_ILP x 32 per step
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
          uint2 row = s_row[r];
          for (uint32_t p0 = pbeg; p0 < pend; p0 += 32 * KL_ILP) {
            uint32_t v[KL_ILP];
#pragma unroll
            for (int j = 0; j < KL_ILP; ++j) {
              const uint32_t p = p0 + 32 * j + lane;
              v[j] = 0xFFFFFFFFu;
              if (p < pend) {
                while (row.x <= p) row = s_row[++r];
                v[j] = fidx[row.y + p];
              }
            }
            // claim, then ONE reservation of queue slots for the step's KL_ILP x 32 entries (32 warps adding to
            // one shared counter per 32 entries showed as the kernel's largest stall)
            unsigned m[KL_ILP];
            uint32_t cnt = 0;
#pragma unroll
            for (int j = 0; j < KL_ILP; ++j) {
              bool isnew = false;
              if (v[j] != 0xFFFFFFFFu) {
                v[j] -= lo;
                const uint32_t bit = 1u << (v[j] & 31);
                if (!(seen[v[j] >> 5] & bit)) isnew = !(atomicOr(&seen[v[j] >> 5], bit) & bit);
              }
              m[j] = __ballot_sync(FULL, isnew);
              cnt += __popc(m[j]);
            }
            if (cnt) {
              uint32_t pos = 0;
              if (lane == 0) pos = atomicAdd(&s_tail, cnt);
              pos = __shfl_sync(FULL, pos, 0);
#pragma unroll
              for (int j = 0; j < KL_ILP; ++j) {
                if ((m[j] >> lane) & 1u) queue[pos + __popc(m[j] & lanemask_lt())] = v[j];
                pos += __popc(m[j]);
              }
            }
          }
        }
        __syncthreads();
      }
      head = tail;
      tail = s_tail;
      nlev = lvl;
      if (tid == 0) s_seg[lvl] = tail;
      __syncthreads();
    }
    const uint32_t total = tail - ns;
    if (tid == 0) {
      const unsigned long long at = atomicAdd(&a.ctr->aux[3], (unsigned long long)total);
      s_base = at;
      a.tmp_start[q] = at;
      a.counts[q] = total;
    }
    if (total) {
      if (a.exclude) {   // excluded sources kept their claim bit during the traversal; they are not results
        for (uint32_t i = tid; i < ns; i += THREADS) {
          const uint32_t v = __ldcg(&queue[i]);
          atomicAnd(&seen[v >> 5], ~(1u << (v & 31)));
        }
      }
      __syncthreads();
      // rank of every quad inside its warp segment, then the segments' bases
      const uint32_t nquads = (wn + 127) >> 7;
      const uint4* seen4 = reinterpret_cast<const uint4*>(seen);
      {
        const uint32_t g0 = (uint32_t)warp << qs_log, g1 = min(nquads, g0 + (1u << qs_log));
        uint32_t run = 0;
        for (uint32_t gq0 = g0; gq0 < g1; gq0 += 32) {
          const uint32_t gq = gq0 + lane;
          uint32_t c = 0;
          if (gq < g1) {
            const uint4 x = seen4[gq];
            c = __popc(x.x) + __popc(x.y) + __popc(x.z) + __popc(x.w);
          }
          const uint32_t inc = warp_incl_scan_u32(c);
          if (gq < g1) rankq[gq] = run + inc - c;
          run += __shfl_sync(FULL, inc, 31);
        }
        if (lane == 0) s_wsum[warp] = run;
      }
      __syncthreads();
      if (warp == 0) {
        const uint32_t x = s_wsum[lane];
        const uint32_t inc = warp_incl_scan_u32(x);
        s_segbase[lane] = inc - x;
      }
      __syncthreads();
      const unsigned long long base = s_base;
      uint32_t segr[KL_SEGR];   // ends of the first levels' queue segments, in registers
#pragma unroll
      for (int l = 1; l <= KL_SEGR; ++l) segr[l - 1] = s_seg[min(l, nlev)];
      // KL_ILP queue entries per thread and step: their (L2) loads are in flight together
      for (uint32_t i0 = 0; i0 < total; i0 += THREADS * KL_ILP) {
        uint32_t vv[KL_ILP];
#pragma unroll
        for (int j = 0; j < KL_ILP; ++j) {
          const uint32_t i = i0 + j * THREADS + tid;
          vv[j] = i < total ? __ldcg(&queue[ns + i]) : 0xFFFFFFFFu;
        }
#pragma unroll
        for (int j = 0; j < KL_ILP; ++j) {
          const uint32_t v = vv[j];
          if (v == 0xFFFFFFFFu) continue;
          const uint32_t p = ns + i0 + j * THREADS + tid, w = v >> 5, gq = w >> 2;
          int lv = 1;   // level segments of the queue: p >= seg[l] puts the entry beyond level l
#pragma unroll
          for (int l = 1; l <= KL_SEGR; ++l) lv += (l < nlev && p >= segr[l - 1]) ? 1 : 0;
          if (lv > KL_SEGR)   // deeper than the registers hold: walk the table
            while (lv < nlev && p >= s_seg[lv]) ++lv;
          const uint4 x = seen4[gq];
          const uint32_t sub = w & 3u;
          uint32_t before = (sub > 0 ? __popc(x.x) : 0) + (sub > 1 ? __popc(x.y) : 0) + (sub > 2 ? __popc(x.z) : 0);
          const uint32_t word = sub == 0 ? x.x : (sub == 1 ? x.y : (sub == 2 ? x.z : x.w));
          before += __popc(word & ((1u << (v & 31)) - 1u));
          const unsigned long long pos = base + s_segbase[gq >> qs_log] + rankq[gq] + before;
          if (pos < a.cap) {
            a.tmp_ids[pos] = lo + v;
            a.tmp_hop[pos] = (uint8_t)lv;
          }
        }
      }
    }
    __syncthreads();
    // reset of the claim bitmap: the whole window with 16-byte shared-memory stores (8 per thread for a
    // million vertices) -- cheaper than ONE round of queue loads from L2, whatever the query touched
    {
      uint4* seen4w = reinterpret_cast<uint4*>(seen);
      for (uint32_t i = tid; i < ((wn + 127) >> 7); i += THREADS) seen4w[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
  }
  te = warp_sum_u64(te);
  rows = warp_sum_u64(rows);
  if (lane == 0 && rows) {
    atomicAdd(&a.ctr->edges, te);
    atomicAdd(&a.ctr->aux[0], te);
    atomicAdd(&a.ctr->rows, rows);
  }
}



}  // namespace dmg
