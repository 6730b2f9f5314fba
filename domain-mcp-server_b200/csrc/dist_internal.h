// Handle bodies of the multi-device layer, shared by dist.cu (replicas, partitioned closure) and
// dist_order.cu (exact analysisOrder on a partitioned graph).
#pragma once
#include <nccl.h>

#include <vector>

#include "common.cuh"

namespace dmg {

// per-rank state of the partitioned closure
struct DistRank {
  cudaStream_t stream = nullptr;
  uint32_t *frontier = nullptr, *visited = nullptr, *next = nullptr, *contrib = nullptr, *gather = nullptr,
           *slice = nullptr, *level = nullptr, *src = nullptr;
  unsigned long long* counters = nullptr;   // [0] edges read, [1] |next|
  size_t src_cap = 0;
  // peer-memory form: the control block of closure_p2p.h
  unsigned char* p2p = nullptr;
  // exact order (dist_order.cu): one allocation holding the rank's depth replica, key and position slices
  unsigned char* ord = nullptr;
};

}  // namespace dmg

struct dmg_dist {
  int n = 0;
  std::vector<int> devices;
  std::vector<dmg_engine*> engines;
  std::vector<ncclComm_t> comms;            // empty until a partitioned graph needs them
  unsigned long long* h_flag = nullptr;     // pinned: |next| of the level
  int peer_state = 0;                       // 0 not probed, 1 every pair has peer access, -1 some pair has not
};

struct dmg_dist_graph {
  dmg_dist* dist = nullptr;
  int partitioned = 0;
  uint32_t V = 0, wr = 0;                   // wr = bitmap words owned per rank; padded width = wr * n
  std::vector<dmg_graph*> graphs;           // one per rank
  std::vector<dmg::DistRank> ranks;
};

namespace dmg {
bool ensure_peer_access(dmg_dist* d);       // dist.cu: peer access between every pair, enabled once
}
