// Exact analysisOrder (PG:390-421) on a vertex-range partitioned graph.
//
// FIFO discovery order == entry points in insertion order, then every BFS level sorted by
// (position of the discovering parent in the order, position of the edge in that parent's dependencies()
// row), the discovering parent being the earliest-ordered parent that has the vertex; unreached vertices
// follow in nodes order (PG:415-418).  The single-GPU form (closure.cu) computes that with a 64-bit
// atomicMin per vertex and an ordered claim -- no sort.  Here the same three passes run per level on every
// rank over the rows of the level's parents it OWNS (an owned vertex's dependencies() row is complete on
// its rank), and everything that belongs to another rank is reached through peer memory:
//     depths    dmg_dist_closure first; every rank then receives all depth slices (one replica each)
//     keys      owner(p) pushes (index of p in its level, edge position) into key[v] on owner(v): a
//               64-bit remote atomicMin, only for neighbours v of the next level
//     count     owner(p) reads key[v] back (remote loads), counts the edges that won, stores the count
//               at p's index of a per-level table on rank 0; rank 0 scans the table
//     place     owner(p) walks the row a third time: a winning edge's vertex gets position
//               level offset + base(parent) + rank among the parent's winners -- stored into the
//               position slice on owner(v) and into the order array on rank 0
// Phases are separate launches on every device with a host-side join in between (no in-kernel
// cross-device synchronisation here; the closure's persistent kernel is where that pays).  Needs peer
// access between all devices of the dmg_dist.
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "closure_p2p.h"
#include "dist_internal.h"

namespace dmg {


struct OrdArgs {
  int rank, n;
  uint32_t V, wr;
  uint32_t v_lo, v_hi;                        // owned vertices
  const uint32_t *ptr, *idx;                  // dependencies() rows (complete for owned vertices)
  const uint32_t* level;                      // this rank's replica of all depths [V]
  unsigned long long* key[P2P_MAX_RANKS];     // key slice of every rank [wr * 32]
  uint32_t* pos[P2P_MAX_RANKS];               // order position slice of every rank [wr * 32]
  uint32_t* cnt;                              // rank 0: winners per parent of the level, by parent index
  const unsigned long long* base;             // rank 0: exclusive scan of cnt
  uint32_t* order;                            // rank 0: the result [V]
  uint32_t lvl;                               // the level being ordered; its parents have depth lvl - 1
  uint32_t off_prev, off_cur;                 // order offsets of level lvl - 1 and of level lvl
};

__device__ __forceinline__ unsigned long long* ord_key(const OrdArgs& a, uint32_t v) {
  const uint32_t owner = (v >> 5) / a.wr;
  return a.key[owner] + (v - owner * a.wr * 32u);
}
__device__ __forceinline__ uint32_t* ord_pos(const OrdArgs& a, uint32_t v) {
  const uint32_t owner = (v >> 5) / a.wr;
  return a.pos[owner] + (v - owner * a.wr * 32u);
}

// PHASE 1 keys, 2 count, 3 place.  One warp per 32 owned vertices; the parents among them are handled one
// after the other, the warp's lanes over the row.
template <int PHASE>
__global__ void __launch_bounds__(256) ord_level_kernel(const OrdArgs a) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t n_words = (a.v_hi - a.v_lo + 31) >> 5;
  const uint32_t* pos_mine = a.pos[a.rank];
  for (uint32_t w = warp; w < n_words; w += nwarps) {
    const uint32_t v0 = a.v_lo + w * 32 + lane;
    unsigned parents = __ballot_sync(FULL, v0 < a.v_hi && a.level[v0] == a.lvl - 1);
    for (; parents; parents &= parents - 1) {
      const uint32_t p = a.v_lo + w * 32 + (__ffs(parents) - 1);
      const uint32_t i = pos_mine[p - a.v_lo] - a.off_prev;   // index of p inside its level
      const uint32_t b = a.ptr[p], n = a.ptr[p + 1] - b;
      if (PHASE == 1) {
        for (uint32_t j = lane; j < n; j += 32) {
          const uint32_t v = a.idx[b + j];
          if (a.level[v] == a.lvl) atomicMin(ord_key(a, v), ((unsigned long long)i << 32) | j);
        }
      } else if (PHASE == 2) {
        uint32_t c = 0;
        for (uint32_t j = lane; j < n; j += 32) {
          const uint32_t v = a.idx[b + j];
          c += a.level[v] == a.lvl && __ldcg(ord_key(a, v)) == (((unsigned long long)i << 32) | j);
        }
        c = warp_sum_u32(c);
        if (lane == 0) a.cnt[i] = c;
      } else {
        uint32_t run = a.off_cur + (uint32_t)__ldcg(&a.base[i]);
        for (uint32_t j0 = 0; j0 < n; j0 += 32) {
          const uint32_t j = j0 + lane;
          bool win = false;
          uint32_t v = 0;
          if (j < n) {
            v = a.idx[b + j];
            win = a.level[v] == a.lvl && __ldcg(ord_key(a, v)) == (((unsigned long long)i << 32) | j);
          }
          const unsigned m = __ballot_sync(FULL, win);
          if (win) {
            const uint32_t at = run + __popc(m & lanemask_lt());
            *ord_pos(a, v) = at;
            a.order[at] = v;
          }
          run += __popc(m);
        }
      }
    }
  }
}

// entry points: positions 0 .. ne-1 in insertion order (already distinct and valid, dmg_graph_create)
__global__ void ord_seed_kernel(const OrdArgs a, const uint32_t* __restrict__ entry, uint32_t ne) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ne) return;
  const uint32_t e = entry[i];
  if (e >= a.v_lo && e < a.v_hi) {
    a.pos[a.rank][e - a.v_lo] = i;
    a.order[i] = e;
  }
}

__global__ void __launch_bounds__(256) ord_count_unreached_kernel(const uint32_t* __restrict__ level, uint32_t v_lo,
                                                                  uint32_t v_hi, unsigned long long* out) {
  unsigned long long c = 0;
  for (uint32_t v = v_lo + blockIdx.x * blockDim.x + threadIdx.x; v < v_hi; v += gridDim.x * blockDim.x)
    c += level[v] == DMG_LEVEL_UNREACHED;
  c = warp_sum_u64(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// orphans of one rank in nodes order (PG:415-418): one CTA walks the slice, a block scan per 1024 vertices
__global__ void __launch_bounds__(1024) ord_tail_kernel(const uint32_t* __restrict__ level, uint32_t v_lo, uint32_t v_hi,
                                                        uint32_t* order, uint32_t at) {
  __shared__ uint32_t s_wsum[33];
  for (uint32_t v0 = v_lo; v0 < v_hi; v0 += 1024) {
    const uint32_t v = v0 + threadIdx.x;
    const uint32_t un = v < v_hi && level[v] == DMG_LEVEL_UNREACHED;
    uint32_t tot;
    const uint32_t off = block_excl_scan<1024>(un, s_wsum, &tot);
    if (un) order[at + off] = v;
    at += tot;
  }
}

}  // namespace dmg

using namespace dmg;

namespace {

inline size_t pad256(size_t x) { return (x + 255) & ~size_t(255); }

struct OrdLayout {   // one device allocation per rank; rank 0's also holds the per-level table and the result
  size_t o_level, o_key, o_pos, o_cnt, o_base, o_order, o_status, o_ctr, bytes;
  OrdLayout(uint32_t V, uint32_t wr, bool rank0) {
    const size_t own = (size_t)wr * 32;
    o_level = 0;
    o_key = pad256(o_level + (size_t)V * 4);
    o_pos = pad256(o_key + own * 8);
    o_cnt = pad256(o_pos + own * 4);
    if (!rank0) {
      o_base = o_order = o_status = o_ctr = bytes = o_cnt;
      return;
    }
    o_base = pad256(o_cnt + (size_t)V * 4);
    o_order = pad256(o_base + ((size_t)V + 1) * 8);
    o_status = pad256(o_order + (size_t)V * 4);
    o_ctr = pad256(o_status + ((size_t)V / (SC_THREADS * SC_ITEMS) + 2) * 8);
    bytes = pad256(o_ctr + sizeof(DevCounters));
  }
};

int32_t join_all(dmg_dist_graph* dg) {
  for (size_t r = 0; r < dg->ranks.size(); ++r) {
    DMG_CUDA(cudaSetDevice(dg->dist->devices[r]));
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaStreamSynchronize(dg->ranks[r].stream));
  }
  return DMG_OK;
}

}  // namespace

extern "C" int32_t dmg_dist_analysis_order(dmg_dist_graph* dg, uint32_t* out_order) {
  if (!dg || !out_order) {
    set_error("dmg_dist_analysis_order: NULL argument");
    return DMG_EINVAL;
  }
  if (!dg->partitioned) {
    set_error("dmg_dist_analysis_order: the graph was created replicated; use dmg_analysis_order on one part");
    return DMG_EINVAL;
  }
  dmg_dist* d = dg->dist;
  const int N = d->n;
  const uint32_t V = dg->V, wr = dg->wr;
  if (V == 0) return DMG_OK;
  if (!ensure_peer_access(d)) {
    set_error("dmg_dist_analysis_order: needs peer access between all devices of the dmg_dist");
    return DMG_EINVAL;
  }
  // ---- depths (also uploads the entry list to every rank: DistRank::src) -----------------------------------
  const std::vector<uint32_t>& entry = dg->graphs[0]->h_entry;   // insertion order, distinct, valid
  std::vector<uint32_t> depth;
  try {
    depth.resize(V);
  } catch (const std::exception&) {
    set_error("dmg_dist_analysis_order: host out of memory");
    return DMG_ENOMEM;
  }
  uint64_t reached = 0;
  DMG_TRY(dmg_dist_closure(dg, DMG_OUT, (uint32_t)entry.size(), entry.data(), depth.data(), &reached, nullptr));
  // ---- per-rank state -------------------------------------------------------------------------------------
  OrdArgs base_args;
  memset(&base_args, 0, sizeof(base_args));
  base_args.n = N;
  base_args.V = V;
  base_args.wr = wr;
  for (int r = 0; r < N; ++r) {
    DistRank& k = dg->ranks[r];
    const OrdLayout L(V, wr, r == 0);
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    if (!k.ord) DMG_CUDA(cudaMalloc((void**)&k.ord, L.bytes));
    DMG_CUDA(cudaMemcpyAsync(k.ord + L.o_level, depth.data(), (size_t)V * 4, cudaMemcpyHostToDevice, k.stream));
    DMG_CUDA(cudaMemsetAsync(k.ord + L.o_key, 0xFF, (size_t)wr * 32 * 8, k.stream));
    base_args.key[r] = reinterpret_cast<unsigned long long*>(k.ord + L.o_key);
    base_args.pos[r] = reinterpret_cast<uint32_t*>(k.ord + L.o_pos);
  }
  const OrdLayout L0(V, wr, true);
  unsigned char* blk0 = dg->ranks[0].ord;
  base_args.cnt = reinterpret_cast<uint32_t*>(blk0 + L0.o_cnt);
  base_args.base = reinterpret_cast<const unsigned long long*>(blk0 + L0.o_base);
  base_args.order = reinterpret_cast<uint32_t*>(blk0 + L0.o_order);
  uint64_t* status0 = reinterpret_cast<uint64_t*>(blk0 + L0.o_status);
  DevCounters* ctr0 = reinterpret_cast<DevCounters*>(blk0 + L0.o_ctr);
  std::vector<OrdArgs> args(N, base_args);
  for (int r = 0; r < N; ++r) {
    dmg_graph* g = dg->graphs[r];
    args[r].rank = r;
    args[r].v_lo = (uint32_t)std::min<uint64_t>(V, (uint64_t)r * wr * 32);
    args[r].v_hi = (uint32_t)std::min<uint64_t>(V, ((uint64_t)r + 1) * wr * 32);
    args[r].ptr = g->d_out_ptr;
    args[r].idx = g->d_out_idx;
    args[r].level = reinterpret_cast<const uint32_t*>(dg->ranks[r].ord + OrdLayout(V, wr, r == 0).o_level);
  }
  DMG_TRY(join_all(dg));
  // ---- level 0: the entry points --------------------------------------------------------------------------
  const uint32_t ne = (uint32_t)entry.size();
  for (int r = 0; r < N && ne; ++r) {
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    ord_seed_kernel<<<(ne + 255) / 256, 256, 0, dg->ranks[r].stream>>>(args[r], dg->ranks[r].src, ne);
  }
  DMG_TRY(join_all(dg));
  // ---- levels ---------------------------------------------------------------------------------------------
  uint64_t off_prev = 0, n_prev = ne, off_cur = ne;
  auto run_phase = [&](int phase) -> int32_t {
    for (int r = 0; r < N; ++r) {
      if (args[r].v_hi <= args[r].v_lo) continue;
      DMG_CUDA(cudaSetDevice(d->devices[r]));
      const unsigned grid = (unsigned)d->engines[r]->sm_count * 8;
      cudaStream_t st = dg->ranks[r].stream;
      if (phase == 1) ord_level_kernel<1><<<grid, 256, 0, st>>>(args[r]);
      else if (phase == 2) ord_level_kernel<2><<<grid, 256, 0, st>>>(args[r]);
      else ord_level_kernel<3><<<grid, 256, 0, st>>>(args[r]);
    }
    return join_all(dg);
  };
  for (uint32_t lvl = 1; n_prev > 0 && lvl < 0x7FFFFFFFu; ++lvl) {
    for (int r = 0; r < N; ++r) {
      args[r].lvl = lvl;
      args[r].off_prev = (uint32_t)off_prev;
      args[r].off_cur = (uint32_t)off_cur;
    }
    DMG_TRY(run_phase(1));
    DMG_TRY(run_phase(2));
    // rank 0: winners per parent -> bases; the total is the size of this level
    DMG_CUDA(cudaSetDevice(d->devices[0]));
    cudaStream_t st0 = dg->ranks[0].stream;
    const uint64_t tiles = (n_prev + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
    DMG_CUDA(cudaMemsetAsync(status0, 0, tiles * 8, st0));
    DMG_CUDA(cudaMemsetAsync(ctr0, 0, sizeof(DevCounters), st0));
    scan_counts_kernel<<<(unsigned)std::min<uint64_t>(tiles, (uint64_t)d->engines[0]->sm_count * 4), SC_THREADS, 0, st0>>>(
        base_args.cnt, n_prev, const_cast<unsigned long long*>(base_args.base), &ctr0->aux[2], status0, ctr0);
    unsigned long long n_cur = 0;
    DMG_CUDA(cudaMemcpyAsync(&n_cur, &ctr0->total, 8, cudaMemcpyDeviceToHost, st0));
    DMG_CUDA(cudaStreamSynchronize(st0));
    if (n_cur == 0) break;
    DMG_TRY(run_phase(3));
    off_prev = off_cur;
    off_cur += n_cur;
    n_prev = n_cur;
  }
  // ---- orphans in nodes order (PG:415-418): ranks own ascending ranges ------------------------------------
  std::vector<unsigned long long> un(N, 0);
  for (int r = 0; r < N; ++r) {
    DistRank& k = dg->ranks[r];
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    DMG_CUDA(cudaMemsetAsync(k.counters, 0, 64, k.stream));
    if (args[r].v_hi > args[r].v_lo)
      ord_count_unreached_kernel<<<d->engines[r]->sm_count * 4, 256, 0, k.stream>>>(args[r].level, args[r].v_lo,
                                                                                   args[r].v_hi, k.counters);
    DMG_CUDA(cudaMemcpyAsync(&un[r], k.counters, 8, cudaMemcpyDeviceToHost, k.stream));
  }
  DMG_TRY(join_all(dg));
  uint64_t at = off_cur;
  for (int r = 0; r < N; ++r) {
    if (un[r]) {
      DMG_CUDA(cudaSetDevice(d->devices[r]));
      ord_tail_kernel<<<1, 1024, 0, dg->ranks[r].stream>>>(args[r].level, args[r].v_lo, args[r].v_hi, base_args.order,
                                                           (uint32_t)at);
    }
    at += un[r];
  }
  DMG_TRY(join_all(dg));
  if (at != V) {
    set_error("dmg_dist_analysis_order: internal error, %llu of %u vertices placed", (unsigned long long)at, V);
    return DMG_ECUDA;
  }
  DMG_CUDA(cudaSetDevice(d->devices[0]));
  DMG_CUDA(cudaMemcpy(out_order, base_args.order, (size_t)V * 4, cudaMemcpyDeviceToHost));
  return DMG_OK;
}
