// Partitioned closure (PG:390-421 reachable set with depths; SURVEY.md 8e, BASELINE config 5) as ONE
// persistent kernel per GPU that does the level's expansion AND its exchange over NVLink peer memory.
//
// The NCCL form of dist.cu runs a level as push/pull kernel -> ncclAllGather -> merge kernel -> 8-byte
// read-back, i.e. per level: N full-width contribution bitmaps received by every rank on the sparse
// (push) levels, three launches per rank and one host round trip.  Here every rank launches one
// cooperative kernel for the whole closure; ranks address each other's buffers directly (one process,
// cudaDeviceEnablePeerAccess) and meet at two flag barriers per level:
//
//   expand   push level (frontier < V/32): a rank walks the rows of the frontier vertices it owns; a
//            neighbour it has not seen is marked in the rank's replica of `visited` (that is the per-rank
//            de-duplication) and its bit is OR-ed into the OWNER's next slice with one remote atomic -- a
//            sparse level moves one 4-byte reduction per newly met vertex instead of N bitmaps;
//            pull level: a rank looks, for every unvisited vertex it owns, for a frontier member in the
//            reverse row (frontier and visited are replicated, so this reads local memory only) and sets
//            the bit in its own next slice
//   barrier A  every push has landed
//   publish  the owner turns its next slice into fresh = next & ~visited, stamps the level, clears the
//            slice, and STORES the fresh word and the new visited word into every rank's replica (only
//            words that change); the rank's fresh count goes to every rank's count table
//   barrier B  every replica is complete; all ranks add up the same counts and take the same branch
//
// A barrier is: __threadfence_system, grid sync, one release store of the epoch into each peer's flag
// row, acquire-poll of the own flag row, grid sync.  Polls give up after `spin_ns` (a peer that never
// launched must not hang the box): the kernel then leaves with status 1 and the host reports DMG_ECUDA.
// Data written by peers is read with ld.global.cg (L2 is the point of coherence for peer stores).
#include <cooperative_groups.h>

#include <algorithm>

#include "common.cuh"
#include "closure_p2p.h"
#include "closure_pull.cuh"

namespace cg = cooperative_groups;

namespace dmg {

__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// all threads of the grid; returns false when a peer did not arrive in time
__device__ __forceinline__ bool peer_barrier(const P2pArgs& a, cg::grid_group& grid, uint32_t epoch) {
  __threadfence_system();
  grid.sync();
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)a.n) {
    const int p = threadIdx.x;
    st_release_sys_u32(&a.flags[p][a.rank], epoch);
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys_u32(&a.flags[a.rank][p]) < epoch) {
      if (global_ns() - t0 > a.spin_ns) {
        a.result->status = 1;
        break;
      }
      __nanosleep(64);
    }
    __threadfence_system();
  }
  grid.sync();
  return __ldcg(&a.result->status) == 0;
}

__global__ void __launch_bounds__(P2P_THREADS) closure_p2p_kernel(const P2pArgs a) {
  cg::grid_group grid = cg::this_grid();
  const int lane = threadIdx.x & 31;
  const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
  const uint32_t warp = gtid >> 5, nwarps = gsz >> 5;
  const uint32_t w_lo = (uint32_t)a.rank * a.wr, w_hi = w_lo + a.wr;   // my words of the padded bitmaps
  uint32_t* frontier = a.frontier[a.rank];
  uint32_t* visited = a.visited[a.rank];
  uint32_t* next_mine = a.next_slice[a.rank];
  unsigned long long te = 0, sent = 0;
  uint32_t epoch = 0;

  // |frontier 0|: every rank holds the same replica, so every rank counts the same number
  {
    unsigned long long c = 0;
    for (uint32_t w = gtid; w < a.wr * (uint32_t)a.n; w += gsz) c += __popc(__ldcg(&frontier[w]));
    c = warp_sum_u64(c);
    if (lane == 0 && c) atomicAdd(&a.acc[0], c);
  }
  grid.sync();
  unsigned long long n_front = __ldcg(&a.acc[0]), total = n_front;
  grid.sync();
  if (gtid == 0) a.acc[0] = 0;
  uint32_t lvl = 1, push_levels = 0, pull_levels = 0;
  bool ok = peer_barrier(a, grid, ++epoch);   // nobody pushes into a slice that is still being set up

  while (ok && n_front > 0 && lvl < 0x7FFFFFFFu) {
    const bool push = n_front * 32 < a.V;
    if (push) {
      for (uint32_t w = w_lo + warp; w < w_hi; w += nwarps) {
        uint32_t bits = __ldcg(&frontier[w]);
        while (bits) {
          const uint32_t u = w * 32 + (__ffs(bits) - 1);
          bits &= bits - 1;
          if (u >= a.V) continue;
          const uint32_t b = a.ptr[u], n = a.ptr[u + 1] - b;
          if (lane == 0) te += n;
          for (uint32_t j = lane; j < n; j += 32) {
            const uint32_t x = a.idx[b + j], xw = x >> 5, bit = 1u << (x & 31);
            if (__ldcg(&visited[xw]) & bit) continue;
            if (xw >= w_lo && xw < w_hi) {
              atomicOr(&next_mine[xw - w_lo], bit);   // mine: the publish step strikes what is visited
            } else if (!(atomicOr(&visited[xw], bit) & bit)) {
              const uint32_t owner = xw / a.wr;
              atomicOr(&a.next_slice[owner][xw - owner * a.wr], bit);
              ++sent;
            }
          }
        }
      }
      ++push_levels;
    } else {
      for (uint32_t w = w_lo + warp; w < w_hi; w += nwarps) {
        uint32_t todo = ~__ldcg(&visited[w]);
        if (w * 32 + 32 > a.V) todo &= a.V > w * 32 ? (0xFFFFFFFFu >> (32 - (a.V - w * 32))) : 0u;
        const uint32_t found = pull_word<true>(a.rptr, a.ridx, frontier, w, todo, a.V, lane, &te);
        if (lane == 0 && found) next_mine[w - w_lo] = found;
      }
      ++pull_levels;
    }
    if (!(ok = peer_barrier(a, grid, ++epoch))) break;

    // publish: one thread per owned word
    unsigned long long fresh_n = 0;
    for (uint32_t i = gtid; i < a.wr; i += gsz) {
      const uint32_t w = w_lo + i;
      const uint32_t nx = __ldcg(&next_mine[i]), vis = __ldcg(&visited[w]), old_f = __ldcg(&frontier[w]);
      const uint32_t fresh = nx & ~vis;
      if (nx) next_mine[i] = 0;
      if (fresh) {
        const uint32_t nv = vis | fresh;
        for (uint32_t bb = fresh; bb; bb &= bb - 1) a.level[i * 32 + (__ffs(bb) - 1)] = lvl;
        for (int p = 0; p < a.n; ++p) a.visited[p][w] = nv;
        fresh_n += __popc(fresh);
        sent += 2 * (a.n - 1);
      }
      if (fresh | old_f) {
        for (int p = 0; p < a.n; ++p) a.frontier[p][w] = fresh;
      }
    }
    fresh_n = warp_sum_u64(fresh_n);
    if (lane == 0 && fresh_n) atomicAdd(&a.acc[0], fresh_n);
    __threadfence_system();
    grid.sync();
    if (blockIdx.x == 0 && threadIdx.x < (unsigned)a.n) {
      const unsigned long long mine = __ldcg(&a.acc[0]);
      a.counts[threadIdx.x][(lvl & 1u) * P2P_MAX_RANKS + a.rank] = mine;
    }
    if (!(ok = peer_barrier(a, grid, ++epoch))) break;
    if (gtid == 0) a.acc[0] = 0;
    n_front = 0;
    for (int r = 0; r < a.n; ++r) n_front += __ldcg(&a.counts[a.rank][(lvl & 1u) * P2P_MAX_RANKS + r]);
    total += n_front;
    ++lvl;
  }
  te = warp_sum_u64(te);
  sent = warp_sum_u64(sent);
  if (lane == 0) {
    if (te) atomicAdd(&a.acc[1], te);
    if (sent) atomicAdd(&a.acc[2], sent);
  }
  grid.sync();
  if (gtid == 0) {
    a.result->reached = total;
    a.result->edges = __ldcg(&a.acc[1]);
    a.result->sent_words = __ldcg(&a.acc[2]);
    a.result->levels = lvl - 1;
    a.result->push_levels = push_levels;
    a.result->pull_levels = pull_levels;
  }
}

int32_t closure_p2p_launch(const P2pArgs& a, int sm_count, cudaStream_t stream) {
  int occ = 0;
  DMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, closure_p2p_kernel, P2P_THREADS, 0));
  if (occ < 1) {
    set_error("closure_p2p_kernel does not fit an SM");
    return DMG_ECUDA;
  }
  // every CTA takes part in 5 grid syncs per level: two resident CTAs per SM hide the memory latency of the
  // expansion without making the syncs slower than they need to be
  const int grid = sm_count * std::min(occ, 2);
  void* params[] = {const_cast<P2pArgs*>(&a)};
  DMG_CUDA(cudaLaunchCooperativeKernel((const void*)closure_p2p_kernel, dim3(grid), dim3(P2P_THREADS), params, 0, stream));
  return DMG_OK;
}

}  // namespace dmg
