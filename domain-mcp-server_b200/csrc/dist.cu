// Multi-GPU inside the C ABI (SURVEY.md 8b "one engine per process, N devices"; 8e): the host this
// library drops into is ONE JVM holding one registry of graphs (GraphService.java:34-38) -- it cannot run
// torchrun.  A dmg_dist owns one engine per device of the box and, for the partitioned closure, one NCCL
// communicator per device created in this process (ncclCommInitAll).  Two modes:
//   replicated   the graph is uploaded to every device; a query batch is cut into contiguous query
//                ranges, one host thread per device runs its range, rows are concatenated in query
//                order.  No collective (C2-C4 of SURVEY.md 8e).
//   partitioned  rank r owns vertices [v_lo, v_hi) and holds the edges with an endpoint there; the
//                closure of PG:390-421 runs level by level (C5 of SURVEY.md 8e).  Where the devices can
//                address each other's memory the whole closure is one persistent kernel per device that
//                exchanges over NVLink peer memory (closure_p2p.cu); otherwise, or with DMG_DIST_NCCL=1,
//                a level is expand (push / pull kernels of closure_part.cu) -> ncclAllGather of frontier
//                bitmaps -> merge, with one 8-byte read-back per level deciding termination.
// NCCL is bound at run time (dlopen of libnccl.so.2) so that a host which never goes multi-GPU needs no
// NCCL, and a process that already loaded one (torch) keeps using that copy.  Failures: DMG_ENCCL.
#include <dlfcn.h>

#include <algorithm>
#include <cstddef>
#include <cstdlib>
#include <thread>
#include <vector>

#include "common.cuh"
#include "closure_p2p.h"
#include "dist_internal.h"

namespace dmg {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok() const { return CommInitAll && CommDestroy && AllGather && GroupStart && GroupEnd && GetErrorString; }
};

static NcclApi* nccl_api() {
  static NcclApi api = [] {
    NcclApi a;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      a.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (a.lib) break;
    }
    if (a.lib) {
      a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(dlsym(a.lib, "ncclCommInitAll"));
      a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(a.lib, "ncclCommDestroy"));
      a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(a.lib, "ncclAllGather"));
      a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(dlsym(a.lib, "ncclGroupStart"));
      a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(dlsym(a.lib, "ncclGroupEnd"));
      a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(a.lib, "ncclGetErrorString"));
    }
    return a;
  }();
  return &api;
}

#define DMG_NCCL(call)                                                                              \
  do {                                                                                              \
    ncclResult_t _r = (call);                                                                       \
    if (_r != ncclSuccess) {                                                                        \
      set_error("NCCL error %d (%s): %s", (int)_r, nccl_api()->GetErrorString(_r), #call);         \
      return DMG_ENCCL;                                                                             \
    }                                                                                               \
  } while (0)

}  // namespace dmg

using namespace dmg;

static void free_rank(DistRank& r) {
  for (void* p : {(void*)r.frontier, (void*)r.visited, (void*)r.next, (void*)r.contrib, (void*)r.gather, (void*)r.slice,
                  (void*)r.level, (void*)r.src, (void*)r.counters, (void*)r.p2p, (void*)r.ord})
    if (p) cudaFree(p);
  if (r.stream) cudaStreamDestroy(r.stream);
  r = DistRank();
}

extern "C" {

int32_t dmg_dist_create(const int32_t* devices, int32_t n, dmg_dist** out) {
  if (!out || n < 1 || n > 64) {
    set_error("dmg_dist_create: need 1..64 devices and an out pointer");
    return DMG_EINVAL;
  }
  *out = nullptr;
  dmg_dist* d = new (std::nothrow) dmg_dist();
  if (!d) return DMG_ENOMEM;
  d->n = n;
  for (int i = 0; i < n; ++i) {
    const int dev = devices ? devices[i] : i;
    for (int j = 0; j < i; ++j)
      if (d->devices[j] == dev) {
        set_error("dmg_dist_create: device %d listed twice", dev);
        dmg_dist_destroy(d);
        return DMG_EINVAL;
      }
    dmg_engine_opts o{dev, 0, 0};
    dmg_engine* e = nullptr;
    const int32_t st = dmg_engine_create(&o, &e);
    if (st != DMG_OK) {
      dmg_dist_destroy(d);
      return st;
    }
    d->devices.push_back(dev);
    d->engines.push_back(e);
  }
  if (cudaMallocHost((void**)&d->h_flag, 64) != cudaSuccess) {
    cudaGetLastError();
    dmg_dist_destroy(d);
    set_error("dmg_dist_create: pinned allocation failed");
    return DMG_ENOMEM;
  }
  *out = d;
  return DMG_OK;
}

void dmg_dist_destroy(dmg_dist* d) {
  if (!d) return;
  for (ncclComm_t c : d->comms)
    if (c) nccl_api()->CommDestroy(c);
  for (dmg_engine* e : d->engines) dmg_engine_destroy(e);
  if (d->h_flag) cudaFreeHost(d->h_flag);
  delete d;
}

int32_t dmg_dist_size(const dmg_dist* d) { return d ? d->n : 0; }

int32_t dmg_dist_engine(dmg_dist* d, int32_t rank, dmg_engine** out) {
  if (!d || !out || rank < 0 || rank >= d->n) {
    set_error("dmg_dist_engine: bad rank");
    return DMG_EINVAL;
  }
  *out = d->engines[rank];
  return DMG_OK;
}

void dmg_dist_graph_release(dmg_dist_graph* dg) {
  if (!dg) return;
  for (size_t r = 0; r < dg->ranks.size(); ++r) {
    cudaSetDevice(dg->dist->devices[r]);
    free_rank(dg->ranks[r]);
  }
  for (dmg_graph* g : dg->graphs) dmg_graph_release(g);
  delete dg;
}

int32_t dmg_dist_graph_part(const dmg_dist_graph* dg, int32_t rank, dmg_graph** out, uint32_t* v_lo, uint32_t* v_hi) {
  if (!dg || rank < 0 || rank >= (int)dg->graphs.size()) {
    set_error("dmg_dist_graph_part: bad rank");
    return DMG_EINVAL;
  }
  if (out) *out = dg->graphs[rank];
  const uint32_t lo = dg->partitioned ? std::min<uint64_t>(dg->V, (uint64_t)rank * dg->wr * 32) : 0;
  const uint32_t hi = dg->partitioned ? std::min<uint64_t>(dg->V, ((uint64_t)rank + 1) * dg->wr * 32) : dg->V;
  if (v_lo) *v_lo = lo;
  if (v_hi) *v_hi = hi;
  return DMG_OK;
}

// Uploads `desc` to every device: whole (partitioned = 0) or, for the closure of a graph too large or too
// hot for one GPU, only the edges with an endpoint in the rank's vertex range (partitioned = 1; names and
// project ranges are not kept: the partitioned form serves dmg_dist_closure only).
int32_t dmg_dist_graph_create(dmg_dist* d, const dmg_graph_desc* desc, int32_t partitioned, dmg_dist_graph** out) {
  if (!d || !desc || !out) {
    set_error("dmg_dist_graph_create: NULL argument");
    return DMG_EINVAL;
  }
  *out = nullptr;
  dmg_dist_graph* dg = new (std::nothrow) dmg_dist_graph();
  if (!dg) return DMG_ENOMEM;
  dg->dist = d;
  dg->partitioned = partitioned ? 1 : 0;
  dg->V = desc->n_vertices;
  const uint32_t words = (desc->n_vertices + 31) / 32;
  dg->wr = std::max<uint32_t>(1, (words + d->n - 1) / d->n);
  dg->graphs.assign(d->n, nullptr);
  std::vector<int32_t> rc(d->n, DMG_OK);
  std::vector<std::string> err(d->n);
  auto build = [&](int r) {
    try {
      dmg_graph_desc mine = *desc;
      std::vector<uint32_t> s, t;
      if (partitioned) {
        const uint64_t lo = (uint64_t)r * dg->wr * 32, hi = lo + (uint64_t)dg->wr * 32;
        for (uint64_t i = 0; i < desc->n_edges; ++i) {
          const uint32_t a = desc->edge_src[i], b = desc->edge_dst[i];
          if ((a >= lo && a < hi) || (b >= lo && b < hi)) {
            s.push_back(a);
            t.push_back(b);
          }
        }
        mine.n_edges = s.size();
        mine.edge_src = s.data();
        mine.edge_dst = t.data();
        mine.name_off = nullptr;
        mine.name_bytes = mine.lower_bytes = nullptr;
        mine.n_parts = 0;
        mine.part_off = nullptr;
      }
      rc[r] = dmg_graph_create(d->engines[r], &mine, &dg->graphs[r]);
      if (rc[r] != DMG_OK) err[r] = dmg_last_error();
    } catch (const std::exception&) {
      rc[r] = DMG_ENOMEM;
      err[r] = "dmg_dist_graph_create: host out of memory";
    }
  };
  {
    std::vector<std::thread> th;
    for (int r = 1; r < d->n; ++r) th.emplace_back(build, r);
    build(0);
    for (auto& x : th) x.join();
  }
  for (int r = 0; r < d->n; ++r)
    if (rc[r] != DMG_OK) {
      set_error("rank %d: %s", r, err[r].c_str());
      const int32_t st = rc[r];
      dmg_dist_graph_release(dg);
      return st;
    }
  *out = dg;
  return DMG_OK;
}

}  // extern "C"

namespace dmg {

__global__ void dist_sources_kernel(const uint32_t* __restrict__ src, uint32_t ns, uint32_t V, uint32_t v_lo, uint32_t v_hi,
                                    uint32_t* frontier, uint32_t* visited, uint32_t* level_slice) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns) return;
  const uint32_t v = src[i];
  if (v >= V) return;   // unknown ids are ignored (PG:194-196)
  atomicOr(&frontier[v >> 5], 1u << (v & 31));
  atomicOr(&visited[v >> 5], 1u << (v & 31));
  if (v >= v_lo && v < v_hi) level_slice[v - v_lo] = 0;
}
__global__ void dist_popc_kernel(const uint32_t* __restrict__ bits, uint32_t words, unsigned long long* out) {
  unsigned long long c = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) c += __popc(bits[i]);
  c = warp_sum_u64(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

static int32_t ensure_comms(dmg_dist* d) {
  if (!d->comms.empty()) return DMG_OK;
  NcclApi* api = nccl_api();
  if (!api->ok()) {
    set_error("NCCL is not available (dlopen libnccl.so.2: %s)", api->lib ? "symbols missing" : "library not found");
    return DMG_ENCCL;
  }
  std::vector<ncclComm_t> comms(d->n, nullptr);
  DMG_NCCL(api->CommInitAll(comms.data(), d->n, d->devices.data()));
  d->comms = comms;
  return DMG_OK;
}

static int32_t ensure_rank_state(dmg_dist_graph* dg) {
  if (!dg->ranks.empty()) return DMG_OK;
  dmg_dist* d = dg->dist;
  const size_t W = (size_t)dg->wr * d->n;
  dg->ranks.assign(d->n, DistRank());
  for (int r = 0; r < d->n; ++r) {
    DistRank& k = dg->ranks[r];
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    DMG_CUDA(cudaStreamCreateWithFlags(&k.stream, cudaStreamNonBlocking));
    for (uint32_t** p : {&k.frontier, &k.visited, &k.next, &k.contrib}) DMG_CUDA(cudaMalloc((void**)p, W * 4 + 16));
    DMG_CUDA(cudaMalloc((void**)&k.gather, W * 4 * d->n + 16));
    DMG_CUDA(cudaMalloc((void**)&k.slice, (size_t)dg->wr * 4 + 16));
    DMG_CUDA(cudaMalloc((void**)&k.level, (size_t)dg->wr * 32 * 4 + 16));
    DMG_CUDA(cudaMalloc((void**)&k.counters, 64));
    DMG_CUDA(cudaMalloc((void**)&k.p2p, P2P_BLOCK));
  }
  return DMG_OK;
}

// Peer access between every pair of the box's devices (NVSwitch: all pairs), enabled once.
bool ensure_peer_access(dmg_dist* d) {
  if (d->peer_state) return d->peer_state > 0;
  d->peer_state = -1;
  if (d->n > P2P_MAX_RANKS) return false;
  for (int r = 0; r < d->n; ++r)
    for (int p = 0; p < d->n; ++p) {
      int can = 0;
      if (p != r && (cudaDeviceCanAccessPeer(&can, d->devices[r], d->devices[p]) != cudaSuccess || !can)) {
        cudaGetLastError();
        return false;
      }
    }
  for (int r = 0; r < d->n; ++r) {
    if (cudaSetDevice(d->devices[r]) != cudaSuccess) return false;
    int coop = 0;
    if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, d->devices[r]) != cudaSuccess || !coop) return false;
    for (int p = 0; p < d->n; ++p) {
      if (p == r) continue;
      const cudaError_t e = cudaDeviceEnablePeerAccess(d->devices[p], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return false;
      }
      cudaGetLastError();
    }
  }
  d->peer_state = 1;
  return true;
}

}  // namespace dmg

extern "C" {

// BFS levels of the whole graph from a source set over `dir` (the reachable set of PG:390-421 with its
// depth), the graph split by vertex range over the devices of `dg`.  out_level has V entries.
int32_t dmg_dist_closure(dmg_dist_graph* dg, int32_t dir, uint32_t n_src, const uint32_t* src, uint32_t* out_level,
                         uint64_t* reached, dmg_dist_closure_stats* stats) {
  if (!dg || !out_level || !reached || (n_src && !src)) {
    set_error("dmg_dist_closure: NULL argument");
    return DMG_EINVAL;
  }
  if (!dg->partitioned) {
    set_error("dmg_dist_closure: the graph was created replicated; use dmg_closure on one part");
    return DMG_EINVAL;
  }
  if (dir != DMG_OUT && dir != DMG_IN) {
    set_error("dmg_dist_closure: dir %d is not DMG_OUT/DMG_IN", dir);
    return DMG_EINVAL;
  }
  dmg_dist* d = dg->dist;
  NcclApi* api = nccl_api();
  const char* force_nccl = getenv("DMG_DIST_NCCL");
  const bool use_p2p = !(force_nccl && force_nccl[0] == '1') && ensure_peer_access(d);
  if (!use_p2p) DMG_TRY(ensure_comms(d));
  DMG_TRY(ensure_rank_state(dg));
  const int N = d->n;
  const uint32_t V = dg->V, wr = dg->wr;
  const size_t W = (size_t)wr * N;
  auto range = [&](int r, uint32_t* lo, uint32_t* hi) {
    *lo = (uint32_t)std::min<uint64_t>(V, (uint64_t)r * wr * 32);
    *hi = (uint32_t)std::min<uint64_t>(V, ((uint64_t)r + 1) * wr * 32);
  };
  // ---- sources -> replicated frontier / visited bitmaps, level 0 stamped by the owners ----------------
  for (int r = 0; r < N; ++r) {
    DistRank& k = dg->ranks[r];
    uint32_t lo, hi;
    range(r, &lo, &hi);
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    if (n_src > k.src_cap) {
      if (k.src) cudaFree(k.src);
      k.src = nullptr;
      DMG_CUDA(cudaMalloc((void**)&k.src, (size_t)n_src * 4));
      k.src_cap = n_src;
    }
    DMG_CUDA(cudaMemsetAsync(k.frontier, 0, W * 4, k.stream));
    DMG_CUDA(cudaMemsetAsync(k.visited, 0, W * 4, k.stream));
    DMG_CUDA(cudaMemsetAsync(k.next, 0, W * 4, k.stream));
    DMG_CUDA(cudaMemsetAsync(k.level, 0xFF, (size_t)wr * 32 * 4, k.stream));
    DMG_CUDA(cudaMemsetAsync(k.counters, 0, 64, k.stream));
    if (n_src) {
      DMG_CUDA(cudaMemcpyAsync(k.src, src, (size_t)n_src * 4, cudaMemcpyHostToDevice, k.stream));
      dist_sources_kernel<<<(n_src + 255) / 256, 256, 0, k.stream>>>(k.src, n_src, V, lo, hi, k.frontier, k.visited, k.level);
    }
    if (r == 0 && !use_p2p) {
      dist_popc_kernel<<<d->engines[0]->sm_count * 2, 256, 0, k.stream>>>(k.frontier, (uint32_t)W, k.counters + 1);
      DMG_CUDA(cudaMemcpyAsync(d->h_flag, k.counters + 1, 8, cudaMemcpyDeviceToHost, k.stream));
    }
    DMG_CUDA(cudaGetLastError());
  }
  uint64_t total = 0, exchanged = 0;
  uint32_t lvl = 1, push_levels = 0, pull_levels = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (use_p2p) {
    // ---- one persistent kernel per device, exchange through peer memory (closure_p2p.cu) ----------------
    P2pArgs base;
    memset(&base, 0, sizeof(base));
    base.n = N;
    base.V = V;
    base.wr = wr;
    base.spin_ns = 2000000000ull;   // a peer that is 2 s late for a barrier is not coming
    for (int r = 0; r < N; ++r) {
      DistRank& k = dg->ranks[r];
      base.frontier[0][r] = k.frontier;
      base.frontier[1][r] = k.next;
      base.visited[r] = k.visited;
      base.next_slice[r] = k.slice;
      base.counts[r] = reinterpret_cast<unsigned long long*>(k.p2p + P2P_O_COUNTS);
      base.flags[r] = reinterpret_cast<uint32_t*>(k.p2p + P2P_O_FLAGS);
    }
    for (int r = 0; r < N; ++r) {   // every buffer a peer may touch is in its start state before anyone launches
      DistRank& k = dg->ranks[r];
      DMG_CUDA(cudaSetDevice(d->devices[r]));
      DMG_CUDA(cudaMemsetAsync(k.slice, 0, (size_t)wr * 4, k.stream));
      DMG_CUDA(cudaMemsetAsync(k.p2p, 0, P2P_BLOCK, k.stream));
    }
    for (int r = 0; r < N; ++r) {
      DMG_CUDA(cudaSetDevice(d->devices[r]));
      DMG_CUDA(cudaStreamSynchronize(dg->ranks[r].stream));
    }
    DMG_CUDA(cudaSetDevice(d->devices[0]));
    DMG_CUDA(cudaEventCreate(&ev0));
    DMG_CUDA(cudaEventCreate(&ev1));
    DMG_CUDA(cudaEventRecord(ev0, dg->ranks[0].stream));
    int32_t launch_rc = DMG_OK;
    for (int r = 0; r < N && launch_rc == DMG_OK; ++r) {
      DistRank& k = dg->ranks[r];
      dmg_graph* g = dg->graphs[r];
      P2pArgs a = base;
      a.rank = r;
      a.ptr = dir == DMG_OUT ? g->d_out_ptr : g->d_in_ptr;
      a.idx = dir == DMG_OUT ? g->d_out_idx : g->d_in_idx;
      a.rptr = dir == DMG_OUT ? g->d_in_ptr : g->d_out_ptr;
      a.ridx = dir == DMG_OUT ? g->d_in_idx : g->d_out_idx;
      a.level = k.level;
      a.acc = reinterpret_cast<unsigned long long*>(k.p2p + P2P_O_ACC);
      a.result = reinterpret_cast<P2pResult*>(k.p2p + P2P_O_RESULT);
      a.bar = reinterpret_cast<uint32_t*>(k.p2p + P2P_O_BAR);
      if (cudaSetDevice(d->devices[r]) != cudaSuccess) launch_rc = DMG_ECUDA;
      else launch_rc = closure_p2p_launch(a, d->engines[r]->sm_count, k.stream);
    }
    // (a rank that failed to launch leaves the others waiting at their first barrier until spin_ns runs out)
    DMG_CUDA(cudaSetDevice(d->devices[0]));
    DMG_CUDA(cudaEventRecord(ev1, dg->ranks[0].stream));
    P2pResult res0{};
    bool timed_out = false;
    for (int r = 0; r < N; ++r) {
      DistRank& k = dg->ranks[r];
      P2pResult res{};
      DMG_CUDA(cudaSetDevice(d->devices[r]));
      DMG_CUDA(cudaStreamSynchronize(k.stream));
      DMG_CUDA(cudaMemcpy(&res, k.p2p + P2P_O_RESULT, sizeof(res), cudaMemcpyDeviceToHost));
      timed_out |= res.status != 0;
      exchanged = std::max<uint64_t>(exchanged, res.sent_words * 4);
      if (r == 0) res0 = res;
    }
    DMG_TRY(launch_rc);
    if (timed_out) {
      cudaEventDestroy(ev0);
      cudaEventDestroy(ev1);
      set_error("dmg_dist_closure: a device did not reach a level barrier within 2 s");
      return DMG_ECUDA;
    }
    total = res0.reached;
    lvl = res0.levels + 1;
    push_levels = res0.push_levels;
    pull_levels = res0.pull_levels;
  } else {
    DMG_CUDA(cudaSetDevice(d->devices[0]));
    DMG_CUDA(cudaStreamSynchronize(dg->ranks[0].stream));
    uint64_t n_front = *d->h_flag;
    total = n_front;
    cudaEventCreate(&ev0);
    cudaEventCreate(&ev1);
    cudaEventRecord(ev0, dg->ranks[0].stream);
    // ---- levels: expand -> allgather -> merge, all ranks issued from this thread -------------------------
    while (n_front > 0) {
      const bool push = n_front < (uint64_t)V / 32;
      for (int r = 0; r < N; ++r) {
        DistRank& k = dg->ranks[r];
        uint32_t lo, hi;
        range(r, &lo, &hi);
        DMG_CUDA(cudaSetDevice(d->devices[r]));
        if (push) DMG_TRY(dmg_closure_push_async(dg->graphs[r], dir, lo, hi, k.frontier, k.contrib, k.counters, k.stream));
        else DMG_TRY(dmg_closure_pull_async(dg->graphs[r], dir, lo, hi, k.frontier, k.visited, k.slice, k.counters, k.stream));
      }
      DMG_NCCL(api->GroupStart());
      for (int r = 0; r < N; ++r) {
        DistRank& k = dg->ranks[r];
        if (push) DMG_NCCL(api->AllGather(k.contrib, k.gather, W, ncclUint32, d->comms[r], k.stream));
        else DMG_NCCL(api->AllGather(k.slice, k.contrib, wr, ncclUint32, d->comms[r], k.stream));
      }
      DMG_NCCL(api->GroupEnd());
      exchanged += push ? W * 4 * N : W * 4;
      for (int r = 0; r < N; ++r) {
        DistRank& k = dg->ranks[r];
        uint32_t lo, hi;
        range(r, &lo, &hi);
        DMG_CUDA(cudaSetDevice(d->devices[r]));
        DMG_TRY(dmg_closure_merge_async(dg->graphs[r], lo, hi, lvl, push ? k.gather : k.contrib, push ? N : 1, (uint32_t)W,
                                        k.visited, k.next, k.level, k.counters + 1, k.stream));
        std::swap(k.frontier, k.next);
      }
      // |next| is identical on every rank: rank 0's copy decides; the others run ahead into the next level
      DMG_CUDA(cudaSetDevice(d->devices[0]));
      DMG_CUDA(cudaMemcpyAsync(d->h_flag, dg->ranks[0].counters + 1, 8, cudaMemcpyDeviceToHost, dg->ranks[0].stream));
      DMG_CUDA(cudaStreamSynchronize(dg->ranks[0].stream));
      n_front = *d->h_flag;
      total += n_front;
      (push ? push_levels : pull_levels)++;
      ++lvl;
      if (lvl > 0x7FFFFFFEu) break;
    }
    cudaEventRecord(ev1, dg->ranks[0].stream);
  }
  // ---- gather the owners' level slices into the caller's array ------------------------------------------
  unsigned long long edges = 0;
  for (int r = 0; r < N; ++r) {
    DistRank& k = dg->ranks[r];
    uint32_t lo, hi;
    range(r, &lo, &hi);
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    if (hi > lo) DMG_CUDA(cudaMemcpyAsync(out_level + lo, k.level, (size_t)(hi - lo) * 4, cudaMemcpyDeviceToHost, k.stream));
  }
  for (int r = 0; r < N; ++r) {
    DistRank& k = dg->ranks[r];
    DMG_CUDA(cudaSetDevice(d->devices[r]));
    unsigned long long e = 0;
    const void* from = use_p2p ? (const void*)(k.p2p + P2P_O_RESULT + offsetof(P2pResult, edges)) : (const void*)k.counters;
    DMG_CUDA(cudaMemcpyAsync(&e, from, 8, cudaMemcpyDeviceToHost, k.stream));
    DMG_CUDA(cudaStreamSynchronize(k.stream));
    edges += e;
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ev0, ev1);
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  *reached = total;
  if (stats) {
    stats->levels = lvl - 1;
    stats->push_levels = push_levels;
    stats->pull_levels = pull_levels;
    stats->transport = use_p2p ? 1u : 0u;
    stats->edges_read = edges;
    stats->exchanged_bytes_per_rank = exchanged;
    stats->device_ms = ms;
  }
  return DMG_OK;
}

}  // extern "C"

// ---- a query batch over N replicas: contiguous query ranges, one host thread per device ---------------
extern "C" int32_t dmg_dist_neighbors_batch(dmg_dist_graph* dg, int32_t mode, int32_t k, int32_t exclude_sources,
                                            uint64_t nq, const uint32_t* q_off, const uint32_t* q_src, uint64_t* out_off,
                                            uint32_t* out_ids, uint8_t* out_hop, uint64_t cap, uint64_t* needed) {
  if (!dg || !needed || !out_off || (nq && !q_off)) {
    set_error("dmg_dist_neighbors_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (dg->partitioned) {
    set_error("dmg_dist_neighbors_batch: needs a replicated graph");
    return DMG_EINVAL;
  }
  const int N = dg->dist->n;
  *needed = 0;
  out_off[0] = 0;
  if (!nq) return DMG_OK;
  // phase 1: every replica sizes its range (row offsets, local to the range); phase 2: with the ranges'
  // bases known, every replica writes its rows straight into the caller's arrays
  std::vector<uint64_t> q_lo(N + 1), need(N, 0), base(N + 1, 0);
  for (int r = 0; r <= N; ++r) q_lo[r] = nq * r / N;
  std::vector<int32_t> rc(N, DMG_OK);
  std::vector<std::string> err(N);
  std::vector<std::vector<uint32_t>> local_off(N);
  // every replica gets its OWN row-offset array: n + 1 entries each, so neighbouring ranges would otherwise
  // share one element of out_off (the end of one range is the start of the next) and race on it
  std::vector<std::vector<uint64_t>> rank_off(N);
  auto run = [&](int r, int phase) {
    const uint64_t a = q_lo[r], b = q_lo[r + 1], n = b - a;
    if (!n) return;
    if (phase == 0) {
      local_off[r].resize(n + 1);
      rank_off[r].assign(n + 1, 0);
      for (uint64_t i = 0; i <= n; ++i) local_off[r][i] = q_off[a + i] - q_off[a];
    }
    uint64_t nd = 0;
    const bool fetch = phase == 1;
    const uint64_t room = fetch ? cap - base[r] : 0;
    const int32_t st = dmg_neighbors_batch(dg->graphs[r], mode, k, exclude_sources, n, local_off[r].data(),
                                           q_src + q_off[a], rank_off[r].data(), fetch ? out_ids + base[r] : nullptr,
                                           fetch ? out_hop + base[r] : nullptr, room, &nd);
    need[r] = nd;
    if (st != DMG_OK && !(phase == 0 && st == DMG_ECAP)) {
      rc[r] = st;
      err[r] = dmg_last_error();
    }
  };
  for (int phase = 0; phase < 2; ++phase) {
    std::vector<std::thread> th;
    for (int r = 1; r < N; ++r) th.emplace_back(run, r, phase);
    run(0, phase);
    for (auto& x : th) x.join();
    for (int r = 0; r < N; ++r)
      if (rc[r] != DMG_OK) {
        set_error("rank %d: %s", r, err[r].c_str());
        return rc[r];
      }
    if (phase == 0) {
      for (int r = 0; r < N; ++r) base[r + 1] = base[r] + need[r];
      *needed = base[N];
      if (base[N] > cap || (base[N] && (!out_ids || !out_hop))) {
        // offsets are still meaningful: rebase them so the caller sees the global layout
        for (int r = 0; r < N; ++r)
          for (uint64_t i = q_lo[r]; i < q_lo[r + 1]; ++i) out_off[i] = rank_off[r][i - q_lo[r]] + base[r];
        out_off[nq] = base[N];
        set_error("dmg_dist_neighbors_batch: capacity %llu < needed %llu", (unsigned long long)cap,
                  (unsigned long long)base[N]);
        return DMG_ECAP;
      }
    }
  }
  for (int r = 0; r < N; ++r)
    for (uint64_t i = q_lo[r]; i < q_lo[r + 1]; ++i) out_off[i] = rank_off[r][i - q_lo[r]] + base[r];
  out_off[nq] = base[N];
  return DMG_OK;
}

// ---- a graph_query batch over N replicas (BASELINE config 4: "query batches sharded across 8 GPUs") ------
// Contiguous op ranges, one host thread per device.  Phase 1 sizes every range (dmg_query_batch without
// output buffers: k-hop groups are counted, not fetched), phase 2 lets every replica write its rows straight
// into the caller's arrays at the range's base.  Catalogs (methods, flags, class types) have to be set on
// every replica (dmg_dist_graph_part hands out the per-device handles).
extern "C" int32_t dmg_dist_query_batch(dmg_dist_graph* dg, uint64_t n_ops, const dmg_query_op* ops,
                                        const uint8_t* text_bytes, uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux,
                                        uint64_t cap, uint64_t* needed) {
  if (!dg || !needed || !out_off || (n_ops && !ops)) {
    set_error("dmg_dist_query_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (dg->partitioned) {
    set_error("dmg_dist_query_batch: needs a replicated graph");
    return DMG_EINVAL;
  }
  const int N = dg->dist->n;
  *needed = 0;
  out_off[0] = 0;
  if (!n_ops) return DMG_OK;
  std::vector<uint64_t> lo(N + 1), need(N, 0), base(N + 1, 0);
  for (int r = 0; r <= N; ++r) lo[r] = n_ops * r / N;
  std::vector<int32_t> rc(N, DMG_OK);
  std::vector<std::string> err(N);
  std::vector<std::vector<uint64_t>> rank_off(N);
  auto run = [&](int r, int phase) {
    const uint64_t a = lo[r], n = lo[r + 1] - a;
    if (!n) return;
    if (phase == 0) rank_off[r].assign(n + 1, 0);
    uint64_t nd = 0;
    const bool fetch = phase == 1;
    const int32_t st = dmg_query_batch(dg->graphs[r], n, ops + a, text_bytes, rank_off[r].data(),
                                       fetch ? out_ids + base[r] : nullptr, fetch ? out_aux + base[r] : nullptr,
                                       fetch ? cap - base[r] : 0, &nd);
    need[r] = nd;
    if (st != DMG_OK && !(phase == 0 && st == DMG_ECAP)) {
      rc[r] = st;
      err[r] = dmg_last_error();
    }
  };
  for (int phase = 0; phase < 2; ++phase) {
    std::vector<std::thread> th;
    for (int r = 1; r < N; ++r) th.emplace_back(run, r, phase);
    run(0, phase);
    for (auto& x : th) x.join();
    for (int r = 0; r < N; ++r)
      if (rc[r] != DMG_OK) {
        set_error("rank %d: %s", r, err[r].c_str());
        return rc[r];
      }
    if (phase == 0) {
      for (int r = 0; r < N; ++r) base[r + 1] = base[r] + need[r];
      *needed = base[N];
      for (int r = 0; r < N; ++r)
        for (uint64_t i = lo[r]; i < lo[r + 1]; ++i) out_off[i] = rank_off[r][i - lo[r]] + base[r];
      out_off[n_ops] = base[N];
      if (base[N] == 0) return DMG_OK;
      if (base[N] > cap || !out_ids || !out_aux) {
        set_error("dmg_dist_query_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)base[N]);
        return DMG_ECAP;
      }
    }
  }
  return DMG_OK;
}
