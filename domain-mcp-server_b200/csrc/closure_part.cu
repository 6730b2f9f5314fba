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
#include "closure_p2p.h"

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

// depth 0, frontier 0 and visited from the source list (unknown ids are ignored, PG:194-196)
__global__ void closure_sources_kernel(const uint32_t* __restrict__ src, uint32_t ns, uint32_t V, uint32_t* frontier,
                                       uint32_t* visited, uint32_t* level) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns) return;
  const uint32_t v = src[i];
  if (v >= V) return;
  atomicOr(&frontier[v >> 5], 1u << (v & 31));
  atomicOr(&visited[v >> 5], 1u << (v & 31));
  level[v] = 0;
}

// rows and adjacency entries of the traversal as PG:390-421 defines them: every reached vertex is
// polled once and its whole row is looked at (the figures the closure reports beside what it READ)
__global__ void __launch_bounds__(CP_THREADS) reached_degree_kernel(const uint32_t* __restrict__ ptr, uint32_t V,
                                                                    const uint32_t* __restrict__ level, DevCounters* ctr) {
  unsigned long long te = 0, rows = 0;
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x)
    if (level[v] != 0xFFFFFFFFu) {
      te += ptr[v + 1] - ptr[v];
      ++rows;
    }
  te = warp_sum_u64(te);
  rows = warp_sum_u64(rows);
  if ((threadIdx.x & 31) == 0 && rows) {
    atomicAdd(&ctr->aux[0], te);
    atomicAdd(&ctr->rows, rows);
  }
}

// Direction-optimising closure on one GPU over bitmaps: the persistent kernel of the partitioned closure
// (closure_p2p.cu) with ONE rank that owns every word -- the whole closure is one cooperative launch, its
// levels meet at grid barriers, nothing returns to the host in between.  Leaves depths in `level`.
int32_t closure_hybrid(CallCtx* c, dmg_graph* g, int dir, uint32_t n_src, const uint32_t* d_src, uint32_t* level,
                       uint64_t* reached, uint64_t* te_out, uint64_t* read_out, uint64_t* rows_out) {
  const uint32_t V = g->info.n_vertices, words = (V + 31) / 32;
  DMG_TRY(c->tmp1.reserve((size_t)words * 4 * 4 + 64));
  DMG_TRY(c->tmp2.reserve(P2P_BLOCK));
  uint32_t* frontier = c->tmp1.as<uint32_t>();
  uint32_t* visited = frontier + words;
  uint32_t* next = visited + words;
  uint32_t* slice = next + words;
  unsigned char* blk = c->tmp2.as<unsigned char>();
  DMG_CUDA(cudaMemsetAsync(frontier, 0, (size_t)words * 16, c->stream));
  DMG_CUDA(cudaMemsetAsync(blk, 0, P2P_BLOCK, c->stream));
  DMG_CUDA(cudaMemsetAsync(level, 0xFF, (size_t)V * 4, c->stream));
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  if (n_src) {
    LaunchScope ls(c, K_CLOSURE_INIT);
    closure_sources_kernel<<<(n_src + 255) / 256, 256, 0, c->stream>>>(d_src, n_src, V, frontier, visited, level);
  }
  P2pArgs a;
  memset(&a, 0, sizeof(a));
  a.rank = 0;
  a.n = 1;
  a.V = V;
  a.wr = words;
  a.ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
  a.idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
  a.rptr = dir == DMG_OUT ? g->d_in_ptr : g->d_out_ptr;
  a.ridx = dir == DMG_OUT ? g->d_in_idx : g->d_out_idx;
  a.frontier[0][0] = frontier;
  a.frontier[1][0] = next;
  a.visited[0] = visited;
  a.next_slice[0] = slice;
  a.counts[0] = reinterpret_cast<unsigned long long*>(blk + P2P_O_COUNTS);
  a.flags[0] = reinterpret_cast<uint32_t*>(blk + P2P_O_FLAGS);
  a.level = level;
  a.acc = reinterpret_cast<unsigned long long*>(blk + P2P_O_ACC);
  a.result = reinterpret_cast<P2pResult*>(blk + P2P_O_RESULT);
  a.bar = reinterpret_cast<uint32_t*>(blk + P2P_O_BAR);
  a.spin_ns = 2000000000ull;
  {
    LaunchScope ls(c, K_CLOSURE_STEP);
    DMG_TRY(closure_p2p_launch(a, g->eng->sm_count, c->stream));
  }
  {
    LaunchScope ls(c, K_ORDER_DEG);
    reached_degree_kernel<<<g->eng->sm_count * 8, CP_THREADS, 0, c->stream>>>(a.ptr, V, level, c->d_counters);
  }
  DMG_CUDA(cudaGetLastError());
  P2pResult res;
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaMemcpyAsync(&res, a.result, sizeof(res), cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  if (res.status) {
    set_error("dmg_closure: the level barrier timed out");
    return DMG_ECUDA;
  }
  *reached = res.reached;
  *te_out = c->h_counters->aux[0];
  *read_out = res.edges;
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
