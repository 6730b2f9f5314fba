// Partitioned closure (PG:390-421 reachable set with depths; SURVEY.md 8e, BASELINE config 5) as ONE
// persistent kernel per GPU that does the level's expansion AND its exchange over NVLink peer memory.
//
// The NCCL form of dist.cu runs a level as push/pull kernel -> ncclAllGather -> merge kernel -> 8-byte
// read-back, i.e. per level: N full-width contribution bitmaps received by every rank on the sparse
// (push) levels, three launches per rank and one host round trip.  Here every rank launches one
// cooperative kernel for the whole closure; ranks address each other's buffers directly (one process,
// cudaDeviceEnablePeerAccess).  Every rank holds replicas of `visited` and of two frontier bitmaps (the
// level being expanded and the one being built); rank r owns the words [r wr, (r+1) wr) of all of them.
//
//   pull level (frontier >= V/32)   a warp looks, for the unvisited vertices of one owned word, for a
//            frontier member in their reverse rows (local reads only) and PUBLISHES the word at once: the
//            fresh bits go into every rank's next-frontier replica, the new visited word into every
//            visited replica, the depths into the owner's level slice.  One barrier ends the level.
//   push level (frontier < V/32)    a rank walks the rows of the frontier vertices it owns; a neighbour
//            it has not seen is marked in the rank's own visited replica (the per-rank de-duplication)
//            and its bit is OR-ed into the OWNER's next slice with one remote atomic -- a sparse level
//            moves one 4-byte reduction per newly met vertex instead of N bitmaps.  Barrier; the owner
//            strikes what was visited, publishes its words as above; barrier.
//
// A barrier costs one pass through a grid-wide counter: every CTA fences and arrives; the CTA that
// arrives last hands the rank's fresh count and the epoch to every peer (release stores into the peers'
// tables), waits for the peers' epochs (acquire polls) and only then releases the other CTAs of its own
// grid.  Polls give up after `spin_ns` (a peer that never launched must not hang the box): the kernel
// then leaves with status 1 and the host reports DMG_ECUDA.  Data written by peers is read with
// ld.global.cg (L2 is the point of coherence for peer stores).
#include <cooperative_groups.h>

#include <algorithm>

#include "common.cuh"
#include "closure_p2p.h"
#include "closure_pull.cuh"

namespace dmg {

__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

constexpr uint32_t BAR_ABORT = 0x80000000u;

// All threads of the grid.  `dirty`: this thread wrote to a peer since the last barrier.  With
// `count_slot` >= 0 the rank's fresh count (a.acc[0], complete once every CTA has arrived) travels with
// the epoch into slot count_slot of every rank's count table.  Returns false when a peer did not arrive.
__device__ __forceinline__ bool peer_barrier(const P2pArgs& a, uint32_t epoch, bool dirty, int count_slot) {
  __shared__ uint32_t s_last;
  if (dirty) __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(&a.bar[0], 1u) + 1u == epoch * gridDim.x;
    if (s_last) __threadfence();
  }
  __syncthreads();
  if (s_last) {
    bool late = false;
    if (threadIdx.x < (unsigned)a.n) {
      const int p = threadIdx.x;
      if (count_slot >= 0) a.counts[p][count_slot * P2P_MAX_RANKS + a.rank] = __ldcg(&a.acc[0]);
      st_release_sys_u32(&a.flags[p][a.rank], epoch);
      const unsigned long long t0 = global_ns();
      while (ld_acquire_sys_u32(&a.flags[a.rank][p]) < epoch) {
        if (global_ns() - t0 > a.spin_ns) {
          late = true;
          break;
        }
        __nanosleep(20);
      }
    }
    late = __syncthreads_or(late);
    if (threadIdx.x == 0) {
      if (late) a.result->status = 1;
      if (count_slot >= 0) a.acc[0] = 0;
      __threadfence_system();
      st_release_gpu_u32(&a.bar[1], late ? (epoch | BAR_ABORT) : epoch);
    }
  } else if (threadIdx.x == 0) {
    while ((ld_acquire_gpu_u32(&a.bar[1]) & ~BAR_ABORT) < epoch) __nanosleep(20);
  }
  __syncthreads();
  return !(__ldcg(&a.bar[1]) & BAR_ABORT);
}

__global__ void __launch_bounds__(P2P_THREADS, P2P_MIN_CTAS) closure_p2p_kernel(const P2pArgs a) {
  const int lane = threadIdx.x & 31;
  const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
  const uint32_t warp = gtid >> 5, nwarps = gsz >> 5;
  const uint32_t w_lo = (uint32_t)a.rank * a.wr, w_hi = w_lo + a.wr;   // my words of the padded bitmaps
  uint32_t* const visited = a.visited[a.rank];
  uint32_t* const next_mine = a.next_slice[a.rank];
  unsigned long long te = 0, sent = 0;
  uint32_t epoch = 0;

  // |frontier 0|: every rank holds the same replica, so every rank counts the same number.  The first
  // barrier also tells that every peer's kernel is running before anybody writes into a peer.
  {
    unsigned long long c = 0;
    const uint32_t* f0 = a.frontier[0][a.rank];
    for (uint32_t w = gtid; w < a.wr * (uint32_t)a.n; w += gsz) c += __popc(__ldcg(&f0[w]));
    c = warp_sum_u64(c);
    if (lane == 0 && c) atomicAdd(&a.acc[3], c);
  }
  bool ok = peer_barrier(a, ++epoch, false, -1);
  unsigned long long n_front = __ldcg(&a.acc[3]), total = n_front;
  uint32_t lvl = 1, push_levels = 0, pull_levels = 0;

  while (ok && n_front > 0 && lvl < 0x7FFFFFFFu) {
    const uint32_t* const frontier = a.frontier[(lvl - 1) & 1][a.rank];   // the level being expanded
    uint32_t* const* const nextf = a.frontier[lvl & 1];                    // every rank's copy of the one being built
    const bool push = n_front * 32 < a.V;
    bool dirty = false;
    unsigned long long fresh_n = 0;
    if (!push) {
      for (uint32_t w = w_lo + warp; w < w_hi; w += nwarps) {
        const uint32_t vis = __ldcg(&visited[w]);
        uint32_t todo = ~vis;
        if (w * 32 + 32 > a.V) todo &= a.V > w * 32 ? (0xFFFFFFFFu >> (32 - (a.V - w * 32))) : 0u;
        const uint32_t found = pull_word<true>(a.rptr, a.ridx, frontier, w, todo, a.V, lane, &te);
        // publish the word: lane p stores into rank p's replicas, the lanes of the fresh bits stamp depths
        if (lane < a.n) {
          nextf[lane][w] = found;
          if (found) a.visited[lane][w] = vis | found;
          dirty = a.n > 1;
        }
        if ((found >> lane) & 1u) a.level[(w - w_lo) * 32 + lane] = lvl;
        if (lane == 0) fresh_n += __popc(found);
      }
      ++pull_levels;
    } else {
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
              dirty = true;
              ++sent;
            }
          }
        }
      }
      ++push_levels;
      if (!(ok = peer_barrier(a, ++epoch, dirty, -1))) break;   // every push has landed
      dirty = false;
      for (uint32_t i = gtid; i < a.wr; i += gsz) {
        const uint32_t w = w_lo + i;
        const uint32_t nx = __ldcg(&next_mine[i]), vis = __ldcg(&visited[w]);
        const uint32_t fresh = nx & ~vis;
        if (nx) next_mine[i] = 0;
        for (int p = 0; p < a.n; ++p) nextf[p][w] = fresh;
        if (fresh) {
          const uint32_t nv = vis | fresh;
          for (uint32_t bb = fresh; bb; bb &= bb - 1) a.level[i * 32 + (__ffs(bb) - 1)] = lvl;
          for (int p = 0; p < a.n; ++p) a.visited[p][w] = nv;
          fresh_n += __popc(fresh);
        }
        dirty = a.n > 1;
      }
    }
    fresh_n = warp_sum_u64(fresh_n);
    if (lane == 0 && fresh_n) atomicAdd(&a.acc[0], fresh_n);
    // every replica complete, every rank's fresh count delivered: all ranks add up the same numbers
    if (!(ok = peer_barrier(a, ++epoch, dirty, (int)(lvl & 1u)))) break;
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
  __threadfence();
  cooperative_groups::this_grid().sync();
  if (gtid == 0) {
    a.result->reached = total;
    a.result->edges = __ldcg(&a.acc[1]);
    a.result->sent_words = __ldcg(&a.acc[2]) + 2ull * (push_levels + pull_levels) * a.wr * (uint32_t)(a.n - 1);
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
  // the cooperative launch guarantees that every CTA is resident, which the barriers rely on; two CTAs
  // per SM hide the memory latency of the expansion
  const int grid = sm_count * std::min(occ, P2P_MIN_CTAS);
  void* params[] = {const_cast<P2pArgs*>(&a)};
  DMG_CUDA(cudaLaunchCooperativeKernel((const void*)closure_p2p_kernel, dim3(grid), dim3(P2P_THREADS), params, 0, stream));
  return DMG_OK;
}

}  // namespace dmg
