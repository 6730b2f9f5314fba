// Bottom-up step of the closure for one 32-vertex word, shared by the single-GPU / NCCL pull kernel
// (closure_part.cu) and the persistent peer-memory kernel (closure_p2p.cu).
#pragma once

namespace dmg {

// Whole warp.  `todo` = the word's vertices that still look for a parent (same value in all lanes).
// Four vertices are in flight, eight lanes each: a group reads its vertex's reverse row in 32-byte
// chunks and stops at the first chunk that holds a frontier member -- on the dense levels that is
// nearly always the first, so a vertex costs one sector instead of its whole row, and the four groups
// keep four independent load chains going.  Returns the mask of vertices that found a parent (all
// lanes); *te (lane-local, summed by the caller over the warp) counts the adjacency entries looked at.
template <bool CG_LOADS>
__device__ __forceinline__ uint32_t pull_word(const uint32_t* __restrict__ rptr, const uint32_t* __restrict__ ridx,
                                              const uint32_t* frontier, uint32_t w, uint32_t todo, uint32_t V, int lane,
                                              unsigned long long* te) {
  const int sub = lane >> 3, sl = lane & 7;
  const unsigned gmask = 0xFFu << (8 * sub);
  uint32_t found = 0;
  while (todo) {
    uint32_t t = todo;
#pragma unroll
    for (int i = 0; i < 3; ++i)
      if (i < sub) t &= t - 1;
#pragma unroll
    for (int i = 0; i < 4; ++i) todo &= todo - 1;
    if (t) {
      const uint32_t bpos = __ffs(t) - 1;
      const uint32_t v = w * 32 + bpos;
      if (v < V) {
        const uint32_t b = rptr[v], n = rptr[v + 1] - b;
        for (uint32_t j0 = 0; j0 < n; j0 += 8) {
          bool hit = false;
          if (j0 + sl < n) {
            const uint32_t u = ridx[b + j0 + sl];
            const uint32_t fw = CG_LOADS ? __ldcg(&frontier[u >> 5]) : frontier[u >> 5];
            hit = (fw >> (u & 31)) & 1u;
          }
          if (sl == 0) *te += min(8u, n - j0);
          if (__ballot_sync(gmask, hit)) {
            found |= 1u << bpos;
            break;
          }
        }
      }
    }
    __syncwarp();
  }
  return __reduce_or_sync(FULL, found);
}

}  // namespace dmg
