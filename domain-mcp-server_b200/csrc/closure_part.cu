// Vertex-range partitioned closure (BASELINE config 5): every rank owns a contiguous
// vertex range, holds the rows of its own vertices, and keeps V-bit frontier / visited bitmaps
// that are replicated on all ranks.  One BFS level is
//     expand (push or pull, this file)  ->  exchange (NCCL allgather, done by the host)  ->
//     merge (this file: OR the gathered bitmaps, strike visited, stamp the owners' levels)
// Push (top-down) is used while the frontier is small: a rank expands the frontier vertices it
// owns and ORs their neighbours into a full-width contribution bitmap; the exchange gathers all
// contribution bitmaps (NCCL has no bitwise-OR reduction).  Pull (bottom-up) is used when the
// frontier is large: a rank looks, for each unvisited vertex it owns, for one frontier member in
// the vertex's reverse row and writes only its own slice of the next bitmap; the exchange
// gathers the slices.
#include <algorithm>
#include <cstddef>

#include "common.cuh"
#include "closure_pull.cuh"

namespace dmg {

constexpr int CP_THREADS = 256;

// one warp per 32-vertex word of the owned range
__global__ void __launch_bounds__(CP_THREADS) part_push_kernel(const uint32_t* __restrict__ ptr,
                                                               const uint32_t* __restrict__ idx, uint32_t w_lo,
                                                               uint32_t w_hi, uint32_t v_lo, uint32_t v_hi,
                                                               const uint32_t* __restrict__ frontier,
                                                               uint32_t* contrib, DevCounters* ctr) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  unsigned long long te = 0;
  for (uint32_t w = w_lo + warp; w < w_hi; w += nwarps) {
    uint32_t bits = frontier[w];
    while (bits) {
      const uint32_t u = w * 32 + (__ffs(bits) - 1);
      bits &= bits - 1;
      if (u < v_lo || u >= v_hi) continue;
      const uint32_t b = ptr[u], n = ptr[u + 1] - b;
      te += n;
      for (uint32_t j = lane; j < n; j += 32) {
        const uint32_t v = idx[b + j], bit = 1u << (v & 31);
        if (!(__ldcg(&contrib[v >> 5]) & bit)) atomicOr(&contrib[v >> 5], bit);
      }
    }
  }
  if (lane == 0 && te) atomicAdd(&ctr->edges, te);
}

__global__ void __launch_bounds__(CP_THREADS) part_pull_kernel(const uint32_t* __restrict__ rptr,
                                                               const uint32_t* __restrict__ ridx, uint32_t w_lo,
                                                               uint32_t w_hi, uint32_t v_lo, uint32_t v_hi,
                                                               const uint32_t* __restrict__ frontier,
                                                               const uint32_t* __restrict__ visited,
                                                               uint32_t* next_slice, DevCounters* ctr) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  unsigned long long te = 0;
  for (uint32_t w = w_lo + warp; w < w_hi; w += nwarps) {
    uint32_t todo = ~visited[w];
    // vertices of the word outside [v_lo, v_hi) belong to a neighbouring rank (ranges start on a multiple
    // of 32, so only the last word can be shared) or lie beyond V
    if (w * 32 + 32 > v_hi) todo &= v_hi > w * 32 ? (0xFFFFFFFFu >> (32 - (v_hi - w * 32))) : 0u;
    const uint32_t found = pull_word<false>(rptr, ridx, frontier, w, todo, v_hi, lane, &te);
    if (lane == 0) next_slice[w - w_lo] = found;
  }
  te = warp_sum_u64(te);
  if (lane == 0 && te) atomicAdd(&ctr->edges, te);
}

// next = (OR over the gathered bitmaps) & ~visited; visited |= next; owners stamp the level
__global__ void __launch_bounds__(CP_THREADS) part_merge_kernel(const uint32_t* __restrict__ gathered,
                                                                uint32_t n_gathered, uint32_t stride_words,
                                                                uint32_t words, uint32_t* visited, uint32_t* next, uint32_t v_lo,
                                                                uint32_t v_hi, uint32_t lvl, uint32_t* level_slice,
                                                                DevCounters* ctr) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t fresh = 0;
  if (w < words) {
    uint32_t x = 0;
    for (uint32_t r = 0; r < n_gathered; ++r) x |= gathered[(size_t)r * stride_words + w];
    const uint32_t vis = visited[w];
    fresh = x & ~vis;
    next[w] = fresh;
    if (fresh) {
      visited[w] = vis | fresh;
      for (uint32_t b = fresh; b; b &= b - 1) {
        const uint32_t v = w * 32 + (__ffs(b) - 1);
        if (v >= v_lo && v < v_hi) level_slice[v - v_lo] = lvl;
      }
    }
  }
  const uint32_t c = warp_sum_u32(__popc(fresh));
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&ctr->total, (unsigned long long)c);
}

// sum of the row lengths of the frontier's vertices (the TE figure of a pull level)
__global__ void __launch_bounds__(CP_THREADS) frontier_degree_kernel(const uint32_t* __restrict__ ptr, uint32_t words,
                                                                     uint32_t V, const uint32_t* __restrict__ frontier,
                                                                     DevCounters* ctr) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long te = 0, rows = 0;
  if (w < words)
    for (uint32_t b = frontier[w]; b; b &= b - 1) {
      const uint32_t v = w * 32 + (__ffs(b) - 1);
      if (v < V) {
        te += ptr[v + 1] - ptr[v];
        ++rows;
      }
    }
  te = warp_sum_u64(te);
  rows = warp_sum_u64(rows);
  if ((threadIdx.x & 31) == 0 && rows) {
    atomicAdd(&ctr->aux[0], te);
    atomicAdd(&ctr->rows, rows);
  }
}

__global__ void source_bits_kernel(const uint32_t* __restrict__ src, uint32_t ns, uint32_t V, uint32_t* bits) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ns && src[i] < V) atomicOr(&bits[src[i] >> 5], 1u << (src[i] & 31));
}

// Direction-optimising closure on one GPU over bitmaps: the same push / pull / merge kernels as
// the partitioned closure with the whole vertex range owned here.  Leaves levels in `level`.
int32_t closure_hybrid(CallCtx* c, dmg_graph* g, int dir, uint32_t n_src, const uint32_t* d_src, uint32_t* level,
                       uint64_t* reached, uint64_t* te_out, uint64_t* read_out, uint64_t* rows_out) {
  const uint32_t V = g->info.n_vertices, words = (V + 31) / 32;
  DMG_TRY(c->tmp1.reserve((size_t)words * 4 * 4 + 64));
  uint32_t* frontier = c->tmp1.as<uint32_t>();
  uint32_t* visited = frontier + words;
  uint32_t* next = visited + words;
  uint32_t* contrib = next + words;
  DMG_CUDA(cudaMemsetAsync(frontier, 0, (size_t)words * 16, c->stream));
  DMG_CUDA(cudaMemsetAsync(level, 0xFF, (size_t)V * 4, c->stream));
  const uint32_t* ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
  const uint32_t* idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
  const uint32_t* rptr = dir == DMG_OUT ? g->d_in_ptr : g->d_out_ptr;
  const uint32_t* ridx = dir == DMG_OUT ? g->d_in_idx : g->d_out_idx;
  const unsigned wgrid = (words + CP_THREADS - 1) / CP_THREADS;
  const unsigned xgrid = (unsigned)std::min<uint64_t>(((uint64_t)words * 32 + CP_THREADS - 1) / CP_THREADS,
                                                      (uint64_t)g->eng->sm_count * 8);
  auto merge = [&](uint32_t lvl, uint64_t* n_new) -> int32_t {
    DMG_CUDA(cudaMemsetAsync(&c->d_counters->total, 0, 8, c->stream));
    {
      LaunchScope ls(c, K_MISC);
      part_merge_kernel<<<wgrid, CP_THREADS, 0, c->stream>>>(contrib, 1, words, words, visited, next, 0, V, lvl, level,
                                                             c->d_counters);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    *n_new = c->h_counters->total;
    return DMG_OK;
  };
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  if (n_src) {
    LaunchScope ls(c, K_CLOSURE_INIT);
    source_bits_kernel<<<(n_src + 255) / 256, 256, 0, c->stream>>>(d_src, n_src, V, contrib);
  }
  uint64_t n_front = 0, total = 0;
  DMG_TRY(merge(0, &n_front));
  total = n_front;
  for (uint32_t lvl = 1; n_front > 0; ++lvl) {
    std::swap(frontier, next);
    {
      LaunchScope ls(c, K_ORDER_DEG);
      frontier_degree_kernel<<<wgrid, CP_THREADS, 0, c->stream>>>(ptr, words, V, frontier, c->d_counters);
    }
    if (n_front * 32 < V) {
      DMG_CUDA(cudaMemsetAsync(contrib, 0, (size_t)words * 4, c->stream));
      LaunchScope ls(c, K_CLOSURE_STEP);
      part_push_kernel<<<xgrid, CP_THREADS, 0, c->stream>>>(ptr, idx, 0, words, 0, V, frontier, contrib, c->d_counters);
    } else {
      LaunchScope ls(c, K_CLOSURE_BOTTOMUP);
      part_pull_kernel<<<xgrid, CP_THREADS, 0, c->stream>>>(rptr, ridx, 0, words, 0, V, frontier, visited, contrib,
                                                            c->d_counters);
    }
    DMG_TRY(merge(lvl, &n_front));
    total += n_front;
  }
  *reached = total;
  *te_out = c->h_counters->aux[0];
  *read_out = c->h_counters->edges;
  *rows_out = c->h_counters->rows;
  return DMG_OK;
}

static int32_t part_args(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi, const char* who) {
  if (!g) {
    set_error("%s: NULL graph", who);
    return DMG_EINVAL;
  }
  if (dir != DMG_OUT && dir != DMG_IN) {
    set_error("%s: dir %d is not DMG_OUT/DMG_IN", who, dir);
    return DMG_EINVAL;
  }
  if (v_lo > v_hi || v_hi > g->info.n_vertices || (v_lo & 31)) {
    set_error("%s: owned range [%u, %u) must lie inside the graph and start on a multiple of 32", who, v_lo, v_hi);
    return DMG_EINVAL;
  }
  return DMG_OK;
}

}  // namespace dmg

using namespace dmg;

extern "C" {

int32_t dmg_closure_push_dev(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi, const uint32_t* d_frontier_bits,
                             uint32_t* d_contrib_bits, uint64_t* edges) {
  DMG_TRY(part_args(g, dir, v_lo, v_hi, "dmg_closure_push_dev"));
  if (!d_frontier_bits || !d_contrib_bits) {
    set_error("dmg_closure_push_dev: NULL bitmap");
    return DMG_EINVAL;
  }
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  const uint32_t words = (g->info.n_vertices + 31) / 32;
  DMG_CUDA(cudaMemsetAsync(d_contrib_bits, 0, (size_t)words * 4, c->stream));
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  const uint32_t w_lo = v_lo / 32, w_hi = (v_hi + 31) / 32;
  if (w_hi > w_lo) {
    const uint32_t* ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
    const uint32_t* idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
    const unsigned grid = (unsigned)std::min<uint64_t>(((uint64_t)(w_hi - w_lo) * 32 + CP_THREADS - 1) / CP_THREADS,
                                                       (uint64_t)g->eng->sm_count * 8);
    LaunchScope ls(c, K_CLOSURE_STEP);
    part_push_kernel<<<grid, CP_THREADS, 0, c->stream>>>(ptr, idx, w_lo, w_hi, v_lo, v_hi, d_frontier_bits,
                                                         d_contrib_bits, c->d_counters);
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_TRY(c->finish());
  if (edges) *edges = c->h_counters->edges;
  c->stats.edges_traversed = c->stats.edges_read = c->h_counters->edges;
  publish_stats(c->stats);
  return DMG_OK;
}

int32_t dmg_closure_pull_dev(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi, const uint32_t* d_frontier_bits,
                             const uint32_t* d_visited_bits, uint32_t* d_next_slice, uint64_t* edges) {
  DMG_TRY(part_args(g, dir, v_lo, v_hi, "dmg_closure_pull_dev"));
  if (!d_frontier_bits || !d_visited_bits || !d_next_slice) {
    set_error("dmg_closure_pull_dev: NULL bitmap");
    return DMG_EINVAL;
  }
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  const uint32_t w_lo = v_lo / 32, w_hi = (v_hi + 31) / 32;
  if (w_hi > w_lo) {
    // closure over dir follows edges u -> v (OUT) : v is reached through its IN row, and vice versa
    const uint32_t* rptr = dir == DMG_OUT ? g->d_in_ptr : g->d_out_ptr;
    const uint32_t* ridx = dir == DMG_OUT ? g->d_in_idx : g->d_out_idx;
    const unsigned grid = (unsigned)std::min<uint64_t>(((uint64_t)(w_hi - w_lo) * 32 + CP_THREADS - 1) / CP_THREADS,
                                                       (uint64_t)g->eng->sm_count * 8);
    LaunchScope ls(c, K_CLOSURE_BOTTOMUP);
    part_pull_kernel<<<grid, CP_THREADS, 0, c->stream>>>(rptr, ridx, w_lo, w_hi, v_lo, v_hi, d_frontier_bits,
                                                         d_visited_bits, d_next_slice, c->d_counters);
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_TRY(c->finish());
  if (edges) *edges = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->edges;
  publish_stats(c->stats);
  return DMG_OK;
}

int32_t dmg_closure_merge_dev(dmg_graph* g, uint32_t v_lo, uint32_t v_hi, uint32_t lvl, const uint32_t* d_gathered,
                              uint32_t n_gathered, uint32_t stride_words, uint32_t* d_visited_bits, uint32_t* d_next_bits,
                              uint32_t* d_level_slice, uint64_t* n_new) {
  DMG_TRY(part_args(g, DMG_OUT, v_lo, v_hi, "dmg_closure_merge_dev"));
  if (!d_gathered || !n_gathered || stride_words < (g->info.n_vertices + 31) / 32 || !d_visited_bits || !d_next_bits || !d_level_slice || !n_new) {
    set_error("dmg_closure_merge_dev: NULL argument");
    return DMG_EINVAL;
  }
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  const uint32_t words = (g->info.n_vertices + 31) / 32;
  if (words) {
    LaunchScope ls(c, K_MISC);
    part_merge_kernel<<<(words + CP_THREADS - 1) / CP_THREADS, CP_THREADS, 0, c->stream>>>(
        d_gathered, n_gathered, stride_words, words, d_visited_bits, d_next_bits, v_lo, v_hi, lvl, d_level_slice, c->d_counters);
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_TRY(c->finish());
  *n_new = c->h_counters->total;
  return DMG_OK;
}


// ---- stream-ordered forms: no context, no synchronisation, no host readback -----------------------
// The host (dmgraph/dist.py) enqueues expand -> ncclAllGather -> merge on ONE stream it owns and
// reads *d_n_new once per level; d_edges (may be NULL) accumulates the adjacency entries read.
int32_t dmg_closure_push_async(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi, const uint32_t* d_frontier_bits,
                               uint32_t* d_contrib_bits, unsigned long long* d_edges, void* cuda_stream) {
  DMG_TRY(part_args(g, dir, v_lo, v_hi, "dmg_closure_push_async"));
  if (!d_frontier_bits || !d_contrib_bits || !d_edges) {
    set_error("dmg_closure_push_async: NULL argument");
    return DMG_EINVAL;
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const uint32_t words = (g->info.n_vertices + 31) / 32;
  DMG_CUDA(cudaMemsetAsync(d_contrib_bits, 0, (size_t)words * 4, st));
  const uint32_t w_lo = v_lo / 32, w_hi = (v_hi + 31) / 32;
  if (w_hi > w_lo) {
    const uint32_t* ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
    const uint32_t* idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
    const unsigned grid = (unsigned)std::min<uint64_t>(((uint64_t)(w_hi - w_lo) * 32 + CP_THREADS - 1) / CP_THREADS,
                                                       (uint64_t)g->eng->sm_count * 8);
    // DevCounters::edges sits 16 bytes into the struct: hand the kernel a base that makes it d_edges
    DevCounters* fake = reinterpret_cast<DevCounters*>(reinterpret_cast<char*>(d_edges) - offsetof(DevCounters, edges));
    part_push_kernel<<<grid, CP_THREADS, 0, st>>>(ptr, idx, w_lo, w_hi, v_lo, v_hi, d_frontier_bits, d_contrib_bits,
                                                  fake);
  }
  DMG_CUDA(cudaGetLastError());
  return DMG_OK;
}

int32_t dmg_closure_pull_async(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi, const uint32_t* d_frontier_bits,
                               const uint32_t* d_visited_bits, uint32_t* d_next_slice, unsigned long long* d_edges,
                               void* cuda_stream) {
  DMG_TRY(part_args(g, dir, v_lo, v_hi, "dmg_closure_pull_async"));
  if (!d_frontier_bits || !d_visited_bits || !d_next_slice || !d_edges) {
    set_error("dmg_closure_pull_async: NULL argument");
    return DMG_EINVAL;
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const uint32_t w_lo = v_lo / 32, w_hi = (v_hi + 31) / 32;
  if (w_hi > w_lo) {
    const uint32_t* rptr = dir == DMG_OUT ? g->d_in_ptr : g->d_out_ptr;
    const uint32_t* ridx = dir == DMG_OUT ? g->d_in_idx : g->d_out_idx;
    const unsigned grid = (unsigned)std::min<uint64_t>(((uint64_t)(w_hi - w_lo) * 32 + CP_THREADS - 1) / CP_THREADS,
                                                       (uint64_t)g->eng->sm_count * 8);
    DevCounters* fake = reinterpret_cast<DevCounters*>(reinterpret_cast<char*>(d_edges) - offsetof(DevCounters, edges));
    part_pull_kernel<<<grid, CP_THREADS, 0, st>>>(rptr, ridx, w_lo, w_hi, v_lo, v_hi, d_frontier_bits, d_visited_bits,
                                                  d_next_slice, fake);
  }
  DMG_CUDA(cudaGetLastError());
  return DMG_OK;
}

int32_t dmg_closure_merge_async(dmg_graph* g, uint32_t v_lo, uint32_t v_hi, uint32_t lvl, const uint32_t* d_gathered,
                                uint32_t n_gathered, uint32_t stride_words, uint32_t* d_visited_bits,
                                uint32_t* d_next_bits, uint32_t* d_level_slice, unsigned long long* d_n_new,
                                void* cuda_stream) {
  DMG_TRY(part_args(g, DMG_OUT, v_lo, v_hi, "dmg_closure_merge_async"));
  if (!d_gathered || !n_gathered || stride_words < (g->info.n_vertices + 31) / 32 || !d_visited_bits || !d_next_bits ||
      !d_level_slice || !d_n_new) {
    set_error("dmg_closure_merge_async: NULL argument");
    return DMG_EINVAL;
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const uint32_t words = (g->info.n_vertices + 31) / 32;
  DMG_CUDA(cudaMemsetAsync(d_n_new, 0, 8, st));
  if (words) {
    DevCounters* fake = reinterpret_cast<DevCounters*>(reinterpret_cast<char*>(d_n_new) - offsetof(DevCounters, total));
    part_merge_kernel<<<(words + CP_THREADS - 1) / CP_THREADS, CP_THREADS, 0, st>>>(
        d_gathered, n_gathered, stride_words, words, d_visited_bits, d_next_bits, v_lo, v_hi, lvl, d_level_slice, fake);
  }
  DMG_CUDA(cudaGetLastError());
  return DMG_OK;
}

}  // extern "C"
