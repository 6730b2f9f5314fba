// Interface between dist.cu (host orchestration) and closure_p2p.cu (the persistent peer-memory kernel).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace dmg {

constexpr int P2P_MAX_RANKS = 16;
constexpr int P2P_THREADS = 512;
constexpr int P2P_MIN_CTAS = 4;      // resident CTAs per SM the kernel is compiled for (pull levels are latency bound)

struct P2pResult {
  unsigned long long reached, edges, sent_words;
  uint32_t levels, push_levels, pull_levels, status;   // status 1: a peer did not reach a barrier in time
};

// Pointers indexed by rank are device addresses on THAT rank's GPU (peer access enabled both ways).
struct P2pArgs {
  int rank, n;
  uint32_t V, wr;                        // wr = bitmap words owned per rank; replicas are wr * n words
  const uint32_t *ptr, *idx;             // rows followed by the closure (owned vertices' rows are complete)
  const uint32_t *rptr, *ridx;           // reverse rows (pull levels)
  uint32_t* frontier[2][P2P_MAX_RANKS];  // replicas of the frontier bitmaps: [level parity][rank]; level 0 is in [0]
  uint32_t* visited[P2P_MAX_RANKS];      // replica of the visited bitmap
  uint32_t* next_slice[P2P_MAX_RANKS];   // [wr] discoveries inside the rank's own range (zero between levels)
  unsigned long long* counts[P2P_MAX_RANKS];   // [2][P2P_MAX_RANKS] fresh vertices per rank, by level parity
  uint32_t* flags[P2P_MAX_RANKS];        // [P2P_MAX_RANKS] barrier epochs, slot r written by rank r
  uint32_t* level;                       // [wr * 32] depth of the owned vertices
  unsigned long long* acc;               // [4] local accumulators (zero at launch): fresh, edges, sent, |frontier 0|
  uint32_t* bar;                         // [2] local grid barrier: arrivals, released epoch (zero at launch)
  P2pResult* result;                     // local (zero at launch)
  unsigned long long spin_ns;
};

// layout of a rank's control block (device memory, zero at launch): counts[2][P2P_MAX_RANKS] u64 |
// flags[P2P_MAX_RANKS] u32 | acc[4] u64 | result | bar[2] u32
constexpr size_t P2P_O_COUNTS = 0, P2P_O_FLAGS = 2 * P2P_MAX_RANKS * 8, P2P_O_ACC = P2P_O_FLAGS + 256,
                 P2P_O_RESULT = P2P_O_ACC + 64, P2P_O_BAR = P2P_O_RESULT + 256, P2P_BLOCK = P2P_O_BAR + 256;

int32_t closure_p2p_launch(const P2pArgs& a, int sm_count, cudaStream_t stream);

}  // namespace dmg
