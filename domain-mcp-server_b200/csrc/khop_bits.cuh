// Bit-parallel batched k-hop for graphs of small projects (included by traverse.cu).
//
// BASELINE config 2 puts 65,536 stack-trace queries on 50 projects of 2,000 classes: ~1,300 queries
// walk the SAME 80,000-entry resolve CSR of one project.  Instead of one traversal per query, a CTA
// takes one project and 64 or 128 of its queries and runs them together (multi-source BFS in the
// sense of Then et al., "The More the Merrier"): every vertex of the project owns one 64/128-bit word
// per set -- bit s = "query slot s" -- so one pass over the project's reverse rows advances all the
// group's traversals at once:
//
//     acc[v]  = OR over u in rev_row(v) of frontier[u]          (one LDS.64/128 + ORs per adjacency entry)
//     new[v]  = acc[v] & ~seen[v];  hop-code planes[v] |= new    (binary-coded hop: bit b of the level)
//
// The reverse rows are read from a per-project sliced-ELL copy built at upload (engine.cu build_ell):
// rows sorted by length, 32 to a slice, lane l of a warp owns row l of the slice and step j holds the
// j-th id of every lane -- coalesced 8-byte loads of four u16 ids per lane, a private OR accumulator
// per lane, no cross-lane step and no atomics in the loop; Zipf hub rows are cut into 256-entry virtual
// rows, so a 1,700-entry hub and thirty 3-entry rows cost the same per entry.  A lane folds its
// accumulator into the planes with shared atomicOr (virtual rows of one vertex meet there), and the
// bits it set first are what it counts as newly found.  No per-query lists, no per-query barriers.
// Level 1 reads the source set, level 2 reads plane 0 (= hop 1 exactly, since nothing else has bit 0
// yet), deeper levels read an explicit frontier buffer.
//
// Result of the traversal kernel: one BIT ROW per query -- NP planes of win_cap/32 words; bit v of
// plane b = bit b of hop(v) (0 = not a neighbour), so k <= 3 needs two planes: 512 bytes per query on
// 2,000-vertex projects against ~10 KB of packed (id, hop) pairs.  The ballot that reads "bit s of 32
// consecutive vertex words" IS the transposition from vertex-major sets to query-major rows.  The
// bit rows (33 MB per config-2 batch, L2-resident) are (a) expanded into the packed, id-ordered rows
// of dmg_neighbors_batch by khop_emit_bits_kernel once scan_counts_kernel has turned the row sizes
// into out_off, or (b) handed to the host as they are (dmg_neighbors_batch_bits): 20x fewer PCIe bytes.
#pragma once

namespace dmg {

constexpr int MB_THREADS = 512;      // threads per project-group CTA
constexpr uint32_t MB_NONE = 0xFFFFFFFFu;
constexpr uint32_t MB_MAX_WINDOW = 4096;   // largest project the shared-memory sets are built for
constexpr int MB_MAX_PLANES = 3;           // hop codes up to 7

struct MbArgs {
  const uint32_t* fptr;      // forward row pointers (TE accounting for the sources)
  // sliced-ELL copy of the reverse rows (what the pull reads)
  const uint32_t* ell_proj_slice;
  const uint32_t* ell_slice_off;
  const uint32_t* ell_vrow;
  const uint2* ell_data;
  uint32_t slots;            // queries per group: 64 or 128 (bits of one set word)
  const uint32_t* part_off;
  uint32_t V, n_parts;
  int k, exclude;
  unsigned long long nq;
  const uint32_t* q_off;
  const uint32_t* q_src;
  uint32_t win_cap;          // largest project rounded up to 128 vertices
  uint32_t WB;               // u32 words per plane of a bit row = win_cap / 32
  // grouping (written by the three small kernels below)
  uint32_t* qpart;           // [nq]  project index of the query's first valid source, MB_NONE if none
  uint32_t* qrank;           // [nq]  arrival order inside the project
  uint32_t* qwin;            // [nq]  (lo, wn)
  uint32_t* part_hist;       // [n_parts] queries per project   (zero before the launch)
  uint32_t* part_gbase;      // [n_parts + 1] first group of each project
  uint32_t* grp_q;           // [max_groups * slots] query of each slot (MB_NONE = empty; preset to 0xFF)
  uint2* grp_win;            // [max_groups] (lo, wn)
  uint32_t* grp_part;        // [max_groups] project index
  uint32_t* n_groups;        // device scalar
  // results
  uint32_t* bits;            // [nq][NP][WB]
  uint32_t* counts;          // [nq] row sizes
  DevCounters* ctr;
};

// ---- grouping: queries -> (project, 64-slot group) ------------------------------------------------
constexpr uint32_t MB_HIST = 2048;   // projects counted in shared memory first
__global__ void __launch_bounds__(256) mb_classify_kernel(const MbArgs a) {
  // Arrival order inside a project: counted per block in shared memory (a few thousand queries hit
  // each of config 2's 50 counters), then one global add per (block, project) fetches the block's base.
  __shared__ uint32_t s_hist[MB_HIST];
  const bool in_smem = a.n_parts <= MB_HIST;
  if (in_smem) {
    for (uint32_t p = threadIdx.x; p < a.n_parts; p += blockDim.x) s_hist[p] = 0;
    __syncthreads();
  }
  const unsigned long long q = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  // The query's graph is the project of its first valid source (CCS:1426-1434, 1558).
  uint32_t part = MB_NONE, lo = 0, hi = 0, local = 0;
  if (q < a.nq) {
    for (uint32_t i = a.q_off[q], e = a.q_off[q + 1]; i < e; ++i) {
      const uint32_t s = a.q_src[i];
      if (s < a.V) {
        uint32_t l = 0, r = a.n_parts;
        while (r - l > 1) {
          const uint32_t m = (l + r) >> 1;
          if (a.part_off[m] <= s) l = m; else r = m;
        }
        part = l;
        lo = a.part_off[l];
        hi = a.part_off[l + 1];
        break;
      }
    }
    a.qpart[q] = part;
    reinterpret_cast<uint2*>(a.qwin)[q] = make_uint2(lo, hi - lo);
    if (part == MB_NONE) a.counts[q] = 0;   // no valid source: empty row, no group
    else local = atomicAdd(in_smem ? &s_hist[part] : &a.part_hist[part], 1u);
  }
  if (in_smem) {
    __syncthreads();
    for (uint32_t p = threadIdx.x; p < a.n_parts; p += blockDim.x) {
      const uint32_t c = s_hist[p];
      if (c) s_hist[p] = atomicAdd(&a.part_hist[p], c);
    }
    __syncthreads();
    if (part != MB_NONE) local += s_hist[part];
  }
  if (part != MB_NONE) a.qrank[q] = local;
}

// One CTA: groups per project = ceil(queries / slots), scanned into part_gbase; every group learns its window.
__global__ void __launch_bounds__(1024) mb_groups_kernel(const MbArgs a) {
  __shared__ uint32_t s_wsum[1024 / 32 + 1];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t p0 = 0; p0 < a.n_parts; p0 += 1024) {
    const uint32_t p = p0 + tid;
    const uint32_t ng = p < a.n_parts ? (a.part_hist[p] + a.slots - 1) / a.slots : 0u;
    uint32_t tot;
    const uint32_t base = s_carry + block_excl_scan<1024>(ng, s_wsum, &tot);
    if (p < a.n_parts) {
      a.part_gbase[p] = base;
      const uint2 w = make_uint2(a.part_off[p], a.part_off[p + 1] - a.part_off[p]);
      for (uint32_t g = 0; g < ng; ++g) {
        a.grp_win[base + g] = w;
        a.grp_part[base + g] = p;
      }
    }
    __syncthreads();
    if (tid == 0) s_carry += tot;
    __syncthreads();
  }
  if (tid == 0) {
    a.part_gbase[a.n_parts] = s_carry;
    *a.n_groups = s_carry;
  }
}

__global__ void __launch_bounds__(256) mb_scatter_kernel(const MbArgs a) {
  const unsigned long long q = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (q >= a.nq) return;
  const uint32_t part = a.qpart[q];
  if (part == MB_NONE) return;
  const uint32_t r = a.qrank[q];
  a.grp_q[(size_t)(a.part_gbase[part] + r / a.slots) * a.slots + (r % a.slots)] = (uint32_t)q;
}

// ---- the traversal ----------------------------------------------------------------------------------
// Shared memory (dynamic): sets of (WC + 8) words of W64 x 8 bytes each -- src, plane[NP] and, for
// k >= 3, two frontier buffers.  Word WC of every set stays zero: the ELL padding id points there.
static inline int msbfs_sets(int np, int k) { return 1 + np + (k >= 3 ? 2 : 0); }
static inline size_t msbfs_smem(uint32_t wc, int np, int k, int w64) {
  return (size_t)msbfs_sets(np, k) * (wc + 8) * 8 * w64;
}

template <int W64> struct SetVec;
template <> struct SetVec<1> { using type = uint2; };
template <> struct SetVec<2> { using type = uint4; };

template <int NP, int W64>
__global__ void __launch_bounds__(MB_THREADS, 2) khop_msbfs_kernel(const MbArgs a) {
  constexpr int THREADS = MB_THREADS, NW = THREADS / 32;
  constexpr int W32 = 2 * W64;          // 32-bit words per set word
  constexpr int SLOTS = 64 * W64;
  using Vec = typename SetVec<W64>::type;
  extern __shared__ __align__(16) unsigned char mb_smem[];
  const uint32_t WC = a.win_cap, WS = WC + 8;   // words per set
  uint32_t* const src = reinterpret_cast<uint32_t*>(mb_smem);   // [WS][W32]
  uint32_t* const plane = src + (size_t)WS * W32;               // [NP][WS][W32]
  uint32_t* const fbuf = plane + (size_t)NP * WS * W32;         // [2][WS][W32], k >= 3 only
  __shared__ uint32_t s_g, s_slice, s_slot_q[SLOTS], s_cnt[SLOTS];
  __shared__ unsigned long long s_te, s_rows, s_reads;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_sets = 1 + NP + (a.k >= 3 ? 2 : 0);
  if (tid == 0) s_te = s_rows = s_reads = 0;

  while (true) {
    __syncthreads();   // the previous group's shared state is no longer read
    if (tid == 0) s_g = atomicAdd(reinterpret_cast<unsigned int*>(&a.ctr->ticket), 1u);
    __syncthreads();
    const uint32_t g = s_g;
    if (g >= *a.n_groups) break;
    const uint2 win = a.grp_win[g];
    const uint32_t lo = win.x, wn = win.y;
    const uint32_t part = a.grp_part[g];
    const uint32_t sl0 = a.ell_proj_slice[part], sl1 = a.ell_proj_slice[part + 1];
    if (tid < SLOTS) {
      s_slot_q[tid] = a.grp_q[(size_t)g * SLOTS + tid];
      s_cnt[tid] = 0;
    }
    {
      uint4* z = reinterpret_cast<uint4*>(mb_smem);
      const uint32_t nz = (uint32_t)((size_t)n_sets * WS * W32 / 4);
      for (uint32_t i = tid; i < nz; i += THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    // L0: bit `slot` of src[v] for every distinct valid source v of the slot's query
    for (int slot = warp; slot < SLOTS; slot += NW) {
      const uint32_t q = s_slot_q[slot];
      if (q == MB_NONE) continue;
      for (uint32_t i = a.q_off[q] + lane, e = a.q_off[q + 1]; i < e; i += 32) {
        const uint32_t s = a.q_src[i];
        if (s >= lo && s - lo < wn) atomicOr(&src[(size_t)(s - lo) * W32 + (slot >> 5)], 1u << (slot & 31));
      }
    }
    __syncthreads();
    unsigned long long te = 0, rows = 0;
    for (uint32_t i = tid; i < wn * W32; i += THREADS) {   // TE of level 1: the sources' forward rows
      const uint32_t f = src[i];
      if (f) {
        const uint32_t v = i / W32, c = __popc(f);
        rows += c;
        te += (unsigned long long)c * (a.fptr[lo + v + 1] - a.fptr[lo + v]);
      }
    }
    uint32_t steps = 0;
    for (int lvl = 1; lvl <= a.k; ++lvl) {
      // frontier of this level; for lvl >= 2 with k >= 3 the new bits are also written to fbuf[lvl & 1]
      const uint32_t* fr = lvl == 1 ? src : (lvl == 2 ? plane : fbuf + (size_t)((lvl - 1) & 1) * WS * W32);
      uint32_t* const fw = (a.k >= 3 && lvl >= 2 && lvl < a.k) ? fbuf + (size_t)(lvl & 1) * WS * W32 : nullptr;
      if (fw) {   // it held the frontier of level lvl - 1: clear before this level's bits arrive
        for (uint32_t i = tid; i < wn * W32; i += THREADS) fw[i] = 0;
      }
      if (tid == 0) s_slice = sl0;
      __syncthreads();
      const Vec* __restrict__ frv = reinterpret_cast<const Vec*>(fr);
      const int b0 = __ffs(lvl) - 1;   // the plane whose atomicOr tells which bits this lane set first
      int any = 0;
      while (true) {
        uint32_t sl = 0;
        if (lane == 0) sl = atomicAdd(&s_slice, 1u);
        sl = __shfl_sync(FULL, sl, 0);
        if (sl >= sl1) break;
        const uint32_t u0 = a.ell_slice_off[sl], u1 = a.ell_slice_off[sl + 1];
        const uint2* __restrict__ d = a.ell_data + u0 + lane;
        uint32_t acc[W32];
#pragma unroll
        for (int w = 0; w < W32; ++w) acc[w] = 0;
        const uint32_t nstep = (u1 - u0) >> 5;
        steps += nstep;
        uint32_t t = 0;
        for (; t + 2 <= nstep; t += 2) {   // 8 ids per lane in flight
          const uint2 x = __ldg(d + (size_t)t * 32), y = __ldg(d + (size_t)(t + 1) * 32);
          const Vec f0 = frv[x.x & 0xFFFFu], f1 = frv[x.x >> 16], f2 = frv[x.y & 0xFFFFu], f3 = frv[x.y >> 16];
          const Vec f4 = frv[y.x & 0xFFFFu], f5 = frv[y.x >> 16], f6 = frv[y.y & 0xFFFFu], f7 = frv[y.y >> 16];
          const uint32_t* p0 = reinterpret_cast<const uint32_t*>(&f0);
          const uint32_t* p1 = reinterpret_cast<const uint32_t*>(&f1);
          const uint32_t* p2 = reinterpret_cast<const uint32_t*>(&f2);
          const uint32_t* p3 = reinterpret_cast<const uint32_t*>(&f3);
          const uint32_t* p4 = reinterpret_cast<const uint32_t*>(&f4);
          const uint32_t* p5 = reinterpret_cast<const uint32_t*>(&f5);
          const uint32_t* p6 = reinterpret_cast<const uint32_t*>(&f6);
          const uint32_t* p7 = reinterpret_cast<const uint32_t*>(&f7);
#pragma unroll
          for (int w = 0; w < W32; ++w) acc[w] |= (p0[w] | p1[w]) | (p2[w] | p3[w]) | (p4[w] | p5[w]) | (p6[w] | p7[w]);
        }
        if (t < nstep) {
          const uint2 x = __ldg(d + (size_t)t * 32);
          const Vec f0 = frv[x.x & 0xFFFFu], f1 = frv[x.x >> 16], f2 = frv[x.y & 0xFFFFu], f3 = frv[x.y >> 16];
          const uint32_t* p0 = reinterpret_cast<const uint32_t*>(&f0);
          const uint32_t* p1 = reinterpret_cast<const uint32_t*>(&f1);
          const uint32_t* p2 = reinterpret_cast<const uint32_t*>(&f2);
          const uint32_t* p3 = reinterpret_cast<const uint32_t*>(&f3);
#pragma unroll
          for (int w = 0; w < W32; ++w) acc[w] |= (p0[w] | p1[w]) | (p2[w] | p3[w]);
        }
        // fold the lane's row into the planes
        const uint32_t vr = a.ell_vrow[(size_t)sl * 32 + lane];
        const uint32_t v = vr & 0xFFFFu, deg = vr >> 16;
        if (v != 0xFFFFu) {
#pragma unroll
          for (int w = 0; w < W32; ++w) {
            if (acc[w] == 0) continue;
            const size_t at = (size_t)v * W32 + w;
            uint32_t sn = a.exclude ? src[at] : 0u;
#pragma unroll
            for (int b = 0; b < NP; ++b) sn |= plane[(size_t)b * WS * W32 + at];
            const uint32_t nw = acc[w] & ~sn;
            if (nw == 0) continue;
            uint32_t fresh = nw;
#pragma unroll
            for (int b = 0; b < NP; ++b) {
              if (!((lvl >> b) & 1)) continue;
              const uint32_t old = atomicOr(&plane[(size_t)b * WS * W32 + at], nw);
              if (b == b0) fresh = nw & ~old;
            }
            if (fresh) {
              any = 1;
              if (lvl < a.k) {   // the next level expands these: TE of SURVEY.md 8d
                const uint32_t c = __popc(fresh);
                rows += c;
                te += (unsigned long long)c * deg;
                if (fw) atomicOr(&fw[at], fresh);
              }
            }
          }
        }
      }
      if (!__syncthreads_or(any)) break;   // no query of the group found anything new
    }
    // ---- bit rows: the sets are vertex-major (word v, bit = slot), a row is query-major (word j, bit =
    // vertex 32 j + i).  A warp loads 32 consecutive vertex words (lane = vertex, one LDS per plane) and
    // transposes each 32 x 32 bit block across its lanes in five shuffle stages, after which lane s holds
    // word j of slot s; two consecutive j are kept so that every lane stores 8 bytes of its row. ----
    const uint32_t WB = a.WB;
    // per-lane constants of the five butterfly stages (j = 16, 8, 4, 2, 1): rotate amount and keep-mask
    uint32_t rot[5], keep[5];
#pragma unroll
    for (int st = 0; st < 5; ++st) {
      const uint32_t j = 16u >> st;
      const uint32_t m = st == 0 ? 0x0000FFFFu : st == 1 ? 0x00FF00FFu : st == 2 ? 0x0F0F0F0Fu : st == 3 ? 0x33333333u : 0x55555555u;
      const bool upper = (lane & j) != 0;
      rot[st] = upper ? 32u - j : j;   // partner's bits move up for the lower lane, down for the upper one
      keep[st] = upper ? ~m : m;       // own bits that stay
    }
    for (uint32_t item = warp; item < WB / 2; item += NW) {   // WB is a multiple of 4 (win_cap of 128)
      uint32_t r[2][W32][NP];
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const uint32_t v = (item * 2 + t) * 32 + lane;
#pragma unroll
        for (int b = 0; b < NP; ++b) {
          const Vec x = reinterpret_cast<const Vec*>(plane + (size_t)b * WS * W32)[v];
          const uint32_t* xw = reinterpret_cast<const uint32_t*>(&x);
#pragma unroll
          for (int w = 0; w < W32; ++w) {
            uint32_t y = xw[w];
#pragma unroll
            for (int st = 0; st < 5; ++st) {
              const uint32_t o = __shfl_xor_sync(FULL, y, 16 >> st);
              y = (y & keep[st]) | (__funnelshift_l(o, o, rot[st]) & ~keep[st]);
            }
            r[t][w][b] = y;
          }
        }
      }
#pragma unroll
      for (int w = 0; w < W32; ++w) {
        const uint32_t q = s_slot_q[32 * w + lane];
        if (q == MB_NONE) continue;
        uint32_t c = 0;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          uint32_t un = 0;
#pragma unroll
          for (int b = 0; b < NP; ++b) un |= r[t][w][b];
          c += __popc(un);
        }
#pragma unroll
        for (int b = 0; b < NP; ++b)
          *reinterpret_cast<uint2*>(a.bits + ((size_t)q * NP + b) * WB + item * 2) = make_uint2(r[0][w][b], r[1][w][b]);
        if (c) atomicAdd(&s_cnt[32 * w + lane], c);
      }
    }
    te = warp_sum_u64(te);
    rows = warp_sum_u64(rows);
    steps = warp_sum_u32(steps);
    if (lane == 0) {
      if (rows) {
        atomicAdd(&s_te, te);
        atomicAdd(&s_rows, rows);
      }
      if (steps) atomicAdd(&s_reads, (unsigned long long)steps * 4);   // ids loaded per lane-step: 4 (padding included)
    }
    __syncthreads();
    if (tid < SLOTS && s_slot_q[tid] != MB_NONE) a.counts[s_slot_q[tid]] = s_cnt[tid];
  }
  if (tid == 0) {
    if (s_rows) {
      atomicAdd(&a.ctr->edges, s_te);
      atomicAdd(&a.ctr->rows, s_rows);
    }
    if (s_reads) atomicAdd(&a.ctr->aux[0], s_reads);
  }
}

// rows of queries that joined no group (no valid source) are all zero in the host-facing bit rows
__global__ void __launch_bounds__(256) mb_zero_orphan_rows_kernel(const uint32_t* __restrict__ qpart, unsigned long long nq,
                                                                  uint32_t* bits, uint32_t row_words) {
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  if (warp >= nq || qpart[warp] != MB_NONE) return;
  for (uint32_t i = lane; i < row_words; i += 32) bits[(size_t)warp * row_words + i] = 0;
}

// 32 x 32 bit transpose across the lanes of a warp: lane i gives row i, receives column i.
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t y, int lane) {
#pragma unroll
  for (int st = 0; st < 5; ++st) {
    const uint32_t j = 16u >> st;
    const uint32_t m = st == 0 ? 0x0000FFFFu : st == 1 ? 0x00FF00FFu : st == 2 ? 0x0F0F0F0Fu : st == 3 ? 0x33333333u : 0x55555555u;
    const bool upper = (lane & j) != 0;
    const uint32_t o = __shfl_xor_sync(FULL, y, j);
    const uint32_t keep = upper ? ~m : m;
    y = (y & keep) | (__funnelshift_l(o, o, upper ? 32u - j : j) & ~keep);
  }
  return y;
}

// ---- bit rows -> packed (id, hop) rows in ascending id: one warp per query ---------------------------
// A warp takes 32 words of every plane (1,024 vertices), transposes them across its lanes so that lane i
// holds, bit by bit, the codes of vertices 32 jj + i, and then walks jj: one ballot ranks the members of
// the 32 vertices of a step, and ids / hops go out as contiguous warp-wide stores.
template <int NP>
__global__ void __launch_bounds__(256, 8) khop_emit_bits_kernel(const uint32_t* __restrict__ bits, uint32_t WB,
                                                                 unsigned long long nq, const uint32_t* __restrict__ qwin,
                                                                 const uint32_t* __restrict__ counts,
                                                                 const unsigned long long* __restrict__ out_off,
                                                                 uint32_t* out_ids, uint8_t* out_hop,
                                                                 unsigned long long cap) {
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
  const unsigned lt = lanemask_lt();
  for (unsigned long long q = warp; q < nq; q += nwarps) {
    const uint32_t cnt = counts[q];
    if (cnt == 0) continue;   // also: queries without a valid source have no bit row
    const unsigned long long row = out_off[q];
    if (row >= cap) continue;
    // entries of this row that fit (all of them unless the caller's capacity ends inside the row)
    const uint32_t room = (uint32_t)min(cap - row, (unsigned long long)cnt);
    const uint32_t lo = qwin[2 * q], wn = qwin[2 * q + 1], nw = (wn + 31) >> 5;
    const uint32_t* __restrict__ br = bits + (size_t)q * NP * WB;
    uint32_t* __restrict__ ids = out_ids + row;
    uint8_t* __restrict__ hops = out_hop + row;
    uint32_t run = 0;
    for (uint32_t j0 = 0; j0 < nw; j0 += 32) {
      uint32_t t[NP];
#pragma unroll
      for (int b = 0; b < NP; ++b) t[b] = warp_transpose32((j0 + lane < nw) ? br[(size_t)b * WB + j0 + lane] : 0u, lane);
      const uint32_t jn = min(32u, nw - j0);
      uint32_t id = lo + j0 * 32 + lane;
#pragma unroll 4
      for (uint32_t jj = 0; jj < jn; ++jj, id += 32) {
        uint32_t code = 0;
#pragma unroll
        for (int b = 0; b < NP; ++b) {
          code |= (t[b] & 1u) << b;
          t[b] >>= 1;
        }
        const uint32_t m = __ballot_sync(FULL, code != 0);
        if (m == FULL) {   // all 32 vertices are members (the common word when the hops cover the project):
          const uint32_t o = run + lane;   // no rank to compute -- POPC issues at a quarter of the ALU rate
          if (o < room) {
            ids[o] = id;
            hops[o] = (uint8_t)code;
          }
          run += 32;
        } else if (m) {
          const uint32_t o = run + __popc(m & lt);
          if (code && o < room) {
            ids[o] = id;
            hops[o] = (uint8_t)code;
          }
          run += __popc(m);
        }
      }
    }
  }
}

}  // namespace dmg
