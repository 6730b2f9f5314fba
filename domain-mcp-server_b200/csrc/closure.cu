// T5 of SURVEY.md 8a: reachability / closure levels and the exact ProjectGraph.analysisOrder()
// sequence (PG:390-421), level-synchronous over the device CSR.
#include <algorithm>

#include "common.cuh"

namespace dmg {

constexpr int CL_THREADS = 256;
constexpr uint32_t HYBRID_MIN_VERTICES = 1u << 15;

struct LevelCounters {
  unsigned int next_count;
  unsigned int pad;
  unsigned long long edges;
};

__global__ void closure_init_kernel(uint32_t* level, uint32_t V, const uint32_t* src, uint32_t ns, uint32_t* frontier,
                                    LevelCounters* lc) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ns) {
    const uint32_t s = src[i];
    if (s < V && atomicCAS(&level[s], DMG_LEVEL_UNREACHED, 0u) == DMG_LEVEL_UNREACHED)
      frontier[atomicAdd(&lc->next_count, 1u)] = s;
  }
}

// Top-down step: one warp per frontier vertex, lanes stride over its row (coalesced 128 B
// loads); a vertex is claimed with atomicCAS on its level word; claimed vertices are appended
// to the next frontier with one atomicAdd per warp iteration (ballot/popc).
__global__ void __launch_bounds__(CL_THREADS) closure_topdown_kernel(const uint32_t* __restrict__ ptr,
                                                                     const uint32_t* __restrict__ idx, uint32_t* level,
                                                                     const uint32_t* __restrict__ frontier, uint32_t nf,
                                                                     uint32_t* next, LevelCounters* lc, uint32_t lvl) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  unsigned long long te = 0;
  for (uint32_t i = warp; i < nf; i += nwarps) {
    const uint32_t u = frontier[i];
    const uint32_t b = ptr[u], n = ptr[u + 1] - b;
    te += n;
    for (uint32_t base = 0; base < n; base += 32) {
      const uint32_t j = base + lane;
      bool isnew = false;
      uint32_t v = 0;
      if (j < n) {
        v = idx[b + j];
        if (__ldcg(&level[v]) == DMG_LEVEL_UNREACHED)
          isnew = atomicCAS(&level[v], DMG_LEVEL_UNREACHED, lvl) == DMG_LEVEL_UNREACHED;
      }
      const unsigned m = __ballot_sync(FULL, isnew);
      if (m) {
        uint32_t pos = 0;
        if (lane == 0) pos = atomicAdd(&lc->next_count, (uint32_t)__popc(m));
        pos = __shfl_sync(FULL, pos, 0);
        if (isnew) next[pos + __popc(m & lanemask_lt())] = v;
      }
    }
  }
  if (lane == 0 && te) atomicAdd(&lc->edges, te);
}

// ---- exact analysisOrder ----------------------------------------------------------------------
// FIFO discovery order == sort each level by (rank of the discovering parent, position of the
// edge in the parent's dependencies() row), where the discovering parent is the lowest-ranked
// parent that has the vertex.  Step 1 computes that minimum key per vertex with a 64-bit
// atomicMin; step 2 walks the parents in rank order (tiles handed out by ticket) and compacts
// the winning edges in order, placing each tile with the look-back allocator -- no sort.
__global__ void __launch_bounds__(CL_THREADS) order_minkey_kernel(const uint32_t* __restrict__ ptr,
                                                                  const uint32_t* __restrict__ idx,
                                                                  const uint32_t* __restrict__ level,
                                                                  const uint32_t* __restrict__ frontier, uint32_t nf,
                                                                  unsigned long long* key) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t i = warp; i < nf; i += nwarps) {
    const uint32_t u = frontier[i];
    const uint32_t b = ptr[u], n = ptr[u + 1] - b;
    for (uint32_t j = lane; j < n; j += 32) {
      const uint32_t v = idx[b + j];
      if (level[v] == DMG_LEVEL_UNREACHED) atomicMin(&key[v], ((unsigned long long)i << 32) | j);
    }
  }
}

constexpr int OC_TILE = 64;   // parents per ticket; 8 warps x 8 parents

__global__ void __launch_bounds__(CL_THREADS) order_claim_kernel(const uint32_t* __restrict__ ptr,
                                                                 const uint32_t* __restrict__ idx, uint32_t* level,
                                                                 const uint32_t* __restrict__ frontier, uint32_t nf,
                                                                 const unsigned long long* __restrict__ key,
                                                                 uint32_t* next, uint32_t* order, uint32_t order_base,
                                                                 uint32_t lvl, uint64_t* status, DevCounters* ctr) {
  __shared__ unsigned long long s_t, s_base;
  __shared__ uint32_t s_cnt[OC_TILE], s_off[OC_TILE];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_tiles = (nf + OC_TILE - 1) / OC_TILE;
  unsigned long long te = 0;   // lane 0 of each warp: adjacency entries of the parents it handled
  while (true) {
    if (tid == 0) s_t = atomicAdd(&ctr->ticket, 1ull);
    __syncthreads();
    const unsigned long long t = s_t;
    if (t >= n_tiles) break;
    // phase A: winners per parent
    for (int p = warp * 8; p < warp * 8 + 8; ++p) {
      const uint32_t i = (uint32_t)t * OC_TILE + p;
      uint32_t c = 0;
      if (i < nf) {
        const uint32_t u = frontier[i];
        const uint32_t b = ptr[u], n = ptr[u + 1] - b;
        te += n;
        for (uint32_t j = lane; j < n; j += 32) {
          const uint32_t v = idx[b + j];
          c += (key[v] == (((unsigned long long)i << 32) | j)) && (__ldcg(&level[v]) == DMG_LEVEL_UNREACHED);
        }
        c = warp_sum_u32(c);
      }
      if (lane == 0) s_cnt[p] = c;
    }
    __syncthreads();
    if (warp == 0) {
      // exclusive scan of the 64 per-parent counts
      const uint32_t a = s_cnt[lane], b2 = s_cnt[lane + 32];
      const uint32_t ia = warp_incl_scan_u32(a);
      const uint32_t tot_a = __shfl_sync(FULL, ia, 31);
      const uint32_t ib = warp_incl_scan_u32(b2);
      s_off[lane] = ia - a;
      s_off[lane + 32] = tot_a + ib - b2;
      const uint32_t total = tot_a + __shfl_sync(FULL, ib, 31);
      const unsigned long long excl = lookback_exclusive(status, t, total);
      if (lane == 0) {
        s_base = excl;
        if (t == n_tiles - 1) ctr->total = excl + total;
      }
    }
    __syncthreads();
    // phase B: ordered placement
    for (int p = warp * 8; p < warp * 8 + 8; ++p) {
      const uint32_t i = (uint32_t)t * OC_TILE + p;
      if (i >= nf) break;
      const uint32_t u = frontier[i];
      const uint32_t b = ptr[u], n = ptr[u + 1] - b;
      uint32_t run = (uint32_t)s_base + s_off[p];
      for (uint32_t base = 0; base < n; base += 32) {
        const uint32_t j = base + lane;
        bool win = false;
        uint32_t v = 0;
        if (j < n) {
          v = idx[b + j];
          win = (key[v] == (((unsigned long long)i << 32) | j)) && (__ldcg(&level[v]) == DMG_LEVEL_UNREACHED);
        }
        const unsigned m = __ballot_sync(FULL, win);
        if (win) {
          const uint32_t pos = run + __popc(m & lanemask_lt());
          next[pos] = v;
          order[order_base + pos] = v;
          level[v] = lvl;
        }
        run += __popc(m);
      }
    }
    __syncthreads();
  }
  if (lane == 0 && te) atomicAdd(&ctr->edges, te);
}

// orphans: every vertex the BFS did not reach, in nodes order (PG:415-418)
__global__ void __launch_bounds__(CL_THREADS) order_tail_kernel(const uint32_t* __restrict__ level, uint32_t V,
                                                                uint32_t* order, uint32_t order_base, uint64_t* status,
                                                                DevCounters* ctr) {
  __shared__ unsigned long long s_t, s_base;
  __shared__ uint32_t s_w[CL_THREADS / 32 + 1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_tiles = (V + CL_THREADS - 1) / CL_THREADS;
  while (true) {
    if (tid == 0) s_t = atomicAdd(&ctr->ticket, 1ull);
    __syncthreads();
    const unsigned long long t = s_t;
    if (t >= n_tiles) break;
    const uint32_t v = (uint32_t)t * CL_THREADS + tid;
    const bool un = v < V && level[v] == DMG_LEVEL_UNREACHED;
    const unsigned m = __ballot_sync(FULL, un);
    if (lane == 0) s_w[warp] = __popc(m);
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = lane < CL_THREADS / 32 ? s_w[lane] : 0;
      const uint32_t iw = warp_incl_scan_u32(w);
      const uint32_t total = __shfl_sync(FULL, iw, 31);
      const unsigned long long excl = lookback_exclusive(status, t, total);
      if (lane < CL_THREADS / 32) s_w[lane] = iw - w;
      if (lane == 0) s_base = excl;
    }
    __syncthreads();
    if (un) order[order_base + (uint32_t)s_base + s_w[warp] + __popc(m & lanemask_lt())] = v;
    __syncthreads();
  }
}

__global__ void order_seed_kernel(const uint32_t* __restrict__ entry, uint32_t ne, uint32_t* level, uint32_t* order,
                                  uint32_t* frontier) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ne) {
    const uint32_t e = entry[i];   // already distinct and valid (dmg_graph_create)
    level[e] = 0;
    order[i] = e;
    frontier[i] = e;
  }
}

}  // namespace dmg

namespace dmg {
int32_t closure_hybrid(CallCtx* c, dmg_graph* g, int dir, uint32_t n_src, const uint32_t* d_src, uint32_t* level,
                       uint64_t* reached, uint64_t* te_out, uint64_t* read_out, uint64_t* rows_out);
}

using namespace dmg;

extern "C" {

int32_t dmg_closure(dmg_graph* g, int32_t dir, uint32_t n_src, const uint32_t* src, uint32_t* out_level,
                    uint64_t* reached) {
  if (!g || !out_level || !reached || (n_src && !src)) {
    set_error("dmg_closure: NULL argument");
    return DMG_EINVAL;
  }
  if (dir != DMG_OUT && dir != DMG_IN) {
    set_error("dmg_closure: dir %d is not DMG_OUT/DMG_IN", dir);
    return DMG_EINVAL;
  }
  const uint32_t V = g->info.n_vertices;
  *reached = 0;
  if (V == 0) return DMG_OK;
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_TRY(c->tmp0.reserve((size_t)V * 4));   // level
  DMG_TRY(c->q_src.reserve((size_t)std::max<uint32_t>(n_src, 1) * 4));
  if (V >= HYBRID_MIN_VERTICES) {
    // large graphs: bitmap frontiers, push while the frontier is small, pull when it is large
    if (n_src) DMG_CUDA(cudaMemcpyAsync(c->q_src.p, src, (size_t)n_src * 4, cudaMemcpyHostToDevice, c->stream));
    c->stats.h2d_bytes += (uint64_t)n_src * 4;
    uint64_t te = 0, rd = 0, rows = 0;
    DMG_TRY(closure_hybrid(c, g, dir, n_src, c->q_src.as<uint32_t>(), c->tmp0.as<uint32_t>(), reached, &te, &rd, &rows));
    DMG_CUDA(cudaMemcpyAsync(out_level, c->tmp0.p, (size_t)V * 4, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += (uint64_t)V * 4;
    c->stats.edges_traversed = te;
    c->stats.edges_read = rd;
    c->stats.rows_expanded = rows;
    c->stats.results = *reached;
    return c->finish();
  }
  DMG_TRY(c->tmp1.reserve((size_t)V * 4));   // frontier A
  DMG_TRY(c->tmp2.reserve((size_t)V * 4));   // frontier B
  uint32_t* level = c->tmp0.as<uint32_t>();
  uint32_t* fa = c->tmp1.as<uint32_t>();
  uint32_t* fb = c->tmp2.as<uint32_t>();
  LevelCounters* lc = reinterpret_cast<LevelCounters*>(c->d_counters);
  LevelCounters* hlc = reinterpret_cast<LevelCounters*>(c->h_counters);
  DMG_CUDA(cudaMemsetAsync(level, 0xFF, (size_t)V * 4, c->stream));
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  if (n_src) {
    DMG_CUDA(cudaMemcpyAsync(c->q_src.p, src, (size_t)n_src * 4, cudaMemcpyHostToDevice, c->stream));
    c->stats.h2d_bytes += (uint64_t)n_src * 4;
    LaunchScope ls(c, K_CLOSURE_INIT);
    closure_init_kernel<<<(n_src + 255) / 256, 256, 0, c->stream>>>(level, V, c->q_src.as<uint32_t>(), n_src, fa, lc);
  }
  DMG_CUDA(cudaMemcpyAsync(hlc, lc, sizeof(LevelCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  uint32_t nf = hlc->next_count;
  uint64_t total = nf, edges = 0, rows = 0;
  const uint32_t* ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
  const uint32_t* idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
  for (uint32_t lvl = 1; nf > 0; ++lvl) {
    DMG_CUDA(cudaMemsetAsync(lc, 0, sizeof(LevelCounters), c->stream));
    const uint64_t warps_needed = nf;
    const unsigned grid = (unsigned)std::min<uint64_t>((warps_needed * 32 + CL_THREADS - 1) / CL_THREADS,
                                                       (uint64_t)g->eng->sm_count * 8);
    {
      LaunchScope ls(c, K_CLOSURE_STEP);
      closure_topdown_kernel<<<grid, CL_THREADS, 0, c->stream>>>(ptr, idx, level, fa, nf, fb, lc, lvl);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(hlc, lc, sizeof(LevelCounters), cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    rows += nf;
    edges += hlc->edges;
    nf = hlc->next_count;
    total += nf;
    std::swap(fa, fb);
  }
  DMG_CUDA(cudaMemcpyAsync(out_level, level, (size_t)V * 4, cudaMemcpyDeviceToHost, c->stream));
  c->stats.d2h_bytes += (uint64_t)V * 4;
  c->stats.edges_traversed = edges;
  c->stats.rows_expanded = rows;
  c->stats.results = total;
  *reached = total;
  return c->finish();
}

int32_t dmg_analysis_order(dmg_graph* g, uint32_t* out_order) {
  if (!g || !out_order) {
    set_error("dmg_analysis_order: NULL argument");
    return DMG_EINVAL;
  }
  const uint32_t V = g->info.n_vertices;
  if (V == 0) return DMG_OK;   // PG:391-393
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_TRY(c->tmp0.reserve((size_t)V * 4));   // level
  DMG_TRY(c->tmp1.reserve((size_t)V * 4));   // frontier A
  DMG_TRY(c->tmp2.reserve((size_t)V * 4));   // frontier B
  DMG_TRY(c->tmp3.reserve((size_t)V * 8));   // min keys
  DMG_TRY(c->out_ids.reserve((size_t)V * 4));  // order
  const size_t max_tiles = ((size_t)V + OC_TILE - 1) / OC_TILE + 1;
  DMG_TRY(c->status.reserve(max_tiles * 8));
  uint32_t* level = c->tmp0.as<uint32_t>();
  uint32_t* fa = c->tmp1.as<uint32_t>();
  uint32_t* fb = c->tmp2.as<uint32_t>();
  unsigned long long* key = c->tmp3.as<unsigned long long>();
  uint32_t* order = c->out_ids.as<uint32_t>();
  DMG_CUDA(cudaMemsetAsync(level, 0xFF, (size_t)V * 4, c->stream));
  DMG_CUDA(cudaMemsetAsync(key, 0xFF, (size_t)V * 8, c->stream));
  uint32_t nf = g->info.n_entry, done = 0;
  uint64_t edges = 0;
  if (nf) {
    LaunchScope ls(c, K_MISC);
    order_seed_kernel<<<(nf + 255) / 256, 256, 0, c->stream>>>(g->d_entry, nf, level, order, fa);
  }
  done = nf;
  for (uint32_t lvl = 1; nf > 0; ++lvl) {
    const unsigned grid = (unsigned)std::min<uint64_t>(((uint64_t)nf * 32 + CL_THREADS - 1) / CL_THREADS,
                                                       (uint64_t)g->eng->sm_count * 8);
    {
      LaunchScope ls(c, K_ORDER_MIN);
      order_minkey_kernel<<<grid, CL_THREADS, 0, c->stream>>>(g->d_out_ptr, g->d_out_idx, level, fa, nf, key);
    }
    const uint32_t n_tiles = (nf + OC_TILE - 1) / OC_TILE;
    DMG_CUDA(cudaMemsetAsync(c->status.p, 0, (size_t)n_tiles * 8, c->stream));
    DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
    {
      LaunchScope ls(c, K_ORDER_CLAIM);
      order_claim_kernel<<<std::min<unsigned>(n_tiles, g->eng->sm_count * 8), CL_THREADS, 0, c->stream>>>(
          g->d_out_ptr, g->d_out_idx, level, fa, nf, key, fb, order, done, lvl, c->status.as<uint64_t>(),
          c->d_counters);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    nf = (uint32_t)c->h_counters->total;
    edges += c->h_counters->edges;   // adjacency entries of this level's parents (TE of SURVEY.md 8d)
    done += nf;
    std::swap(fa, fb);
  }
  if (done < V) {
    const uint32_t n_tiles = (V + CL_THREADS - 1) / CL_THREADS;
    DMG_TRY(c->status.reserve((size_t)n_tiles * 8));
    DMG_CUDA(cudaMemsetAsync(c->status.p, 0, (size_t)n_tiles * 8, c->stream));
    DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
    LaunchScope ls(c, K_ORDER_TAIL);
    order_tail_kernel<<<std::min<unsigned>(n_tiles, g->eng->sm_count * 8), CL_THREADS, 0, c->stream>>>(
        level, V, order, done, c->status.as<uint64_t>(), c->d_counters);
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(out_order, order, (size_t)V * 4, cudaMemcpyDeviceToHost, c->stream));
  c->stats.d2h_bytes += (uint64_t)V * 4;
  c->stats.edges_traversed = edges;
  c->stats.results = V;
  return c->finish();
}

}  // extern "C"
