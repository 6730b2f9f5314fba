// Internal declarations shared by the .cu files of libdmgraph.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dmgraph.h"

namespace dmg {

// FNV-1a over the bytes, then a finaliser; the same function builds the tables on the host and
// probes them on the device
#ifdef __CUDACC__
__host__ __device__
#endif
inline uint32_t name_hash(const uint8_t* p, uint32_t n) {
  unsigned long long h = 0xCBF29CE484222325ull;
  for (uint32_t i = 0; i < n; ++i) h = (h ^ p[i]) * 0x100000001B3ull;
  h ^= h >> 32;
  h *= 0xD6E8FEB86659FD93ull;
  return (uint32_t)(h >> 32);
}

// ---------------------------------------------------------------------------------------------
// errors: thread-local message, status codes; nothing is ever printed (MCP stdio hygiene)
// ---------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int32_t cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define DMG_CUDA(call)                                                         \
  do {                                                                         \
    cudaError_t _e = (call);                                                   \
    if (_e != cudaSuccess) return ::dmg::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define DMG_TRY(expr)                 \
  do {                                \
    int32_t _s = (expr);              \
    if (_s != DMG_OK) return _s;      \
  } while (0)

// ---------------------------------------------------------------------------------------------
// growable device buffer owned by a call context
// ---------------------------------------------------------------------------------------------
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int32_t reserve(size_t bytes);   // grows (never shrinks); contents are NOT preserved
  void release();
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

enum KernelId : int {
  K_KHOP_SMALL = 0,
  K_KHOP_GLOBAL,
  K_KHOP_EMIT,
  K_SCAN_COUNTS,
  K_ADJ_GATHER,
  K_ENTRY_FLAGS,
  K_CLOSURE_INIT,
  K_CLOSURE_STEP,
  K_CLOSURE_BOTTOMUP,
  K_ORDER_DEG,
  K_ORDER_MIN,
  K_ORDER_CLAIM,
  K_ORDER_TAIL,
  K_SEARCH_SCAN,
  K_SEARCH_FIRST,
  K_SEARCH_EMIT,
  K_KHOP_WINDOWS,
  K_QUERY_ROWS,
  K_MISC,
  K_KHOP1_ROWS,
  K_KHOP_LARGE,
  K_SEARCH_REPLICATE,
  K_COUNT
};
extern const char* const kKernelNames[K_COUNT];

struct Engine;

// Counters a kernel family reports back (device memory, zeroed per call, read back pinned).
struct DevCounters {
  unsigned long long ticket;     // dynamic work distribution
  unsigned long long total;      // total entries (== out_off[n])
  unsigned long long edges;      // adjacency entries read
  unsigned long long rows;       // rows expanded
  unsigned long long aux[4];
};

// One in-flight call: its own stream, workspace and event pool.  Handed out by the engine so
// that any number of host threads can query one graph concurrently (SURVEY 8b "Threading").
struct CallCtx {
  Engine* eng = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  DevBuf status, scratch, q_off, q_src, q_bytes, out_off, out_ids, out_hop, tmp0, tmp1, tmp2, tmp3;
  DevCounters* d_counters = nullptr;
  DevCounters* h_counters = nullptr;   // pinned
  void* h_pin = nullptr;               // pinned staging of the host-pointer calls that expand on the host
  size_t h_pin_cap = 0;
  int32_t reserve_pinned(size_t bytes);   // grows, never shrinks; contents are not preserved
  dmg_call_stats stats{};
  struct ProfEv { int kid; cudaEvent_t a, b; };
  std::vector<ProfEv> prof;            // events recorded during this call
  std::vector<cudaEvent_t> ev_pool;
  int32_t init(Engine* e);
  void destroy();
  cudaEvent_t get_event();
  int32_t begin();                     // zero stats, record ev_begin
  int32_t finish();                    // record ev_end, sync, harvest profile events
};

struct Engine {
  int device = 0;
  int sm_count = 148;
  size_t smem_optin = 0;
  size_t scratch_budget = 4ull << 30;
  size_t l2_bytes = 0;
  std::mutex mu;
  std::vector<CallCtx*> free_ctx, all_ctx;
  std::atomic<int> profile{0};
  std::mutex prof_mu;
  uint64_t prof_launches[K_COUNT] = {0};
  double prof_ms[K_COUNT] = {0};
  void* flush_buf = nullptr;
  size_t flush_bytes = 0;
  CallCtx* acquire();
  void release(CallCtx* c);
  // Device blocks handed back by finished dmg_query_batch calls, kept for the next call: a k-hop group's
  // rows are gigabytes, and a cudaMalloc + cudaFree pair per group and call costs more than the traversal
  // that fills them.  Bounded by pool_limit bytes (a quarter of the device) and a handful of blocks.
  std::mutex pool_mu;
  std::vector<std::pair<void*, size_t>> pool;
  size_t pool_bytes = 0, pool_limit = 0;
  void* pool_take(size_t bytes, size_t* got);   // nullptr: device out of memory even after the pool was emptied
  void pool_give(void* p, size_t bytes);
  void pool_clear();
};

struct CtxGuard {
  Engine* e;
  CallCtx* c;
  explicit CtxGuard(Engine* en) : e(en), c(en->acquire()) {}
  // A call that fails half way (a reserve() that runs out of memory after kernels or copies were
  // queued) returns through here with work still in flight on the context's stream: wait for it
  // before another thread may acquire the context and reallocate or overwrite its buffers.  On
  // the normal path finish() has already synchronised and the query returns at once.
  ~CtxGuard() {
    if (!c) return;
    if (cudaStreamQuery(c->stream) != cudaSuccess) {
      cudaStreamSynchronize(c->stream);
      cudaGetLastError();
    }
    e->release(c);
  }
};

void publish_stats(const dmg_call_stats& s);   // thread-local "last call" record

// bits_decode.cc: rows [q_begin, q_end) of a bit-row result into packed (id, hop) rows on the host
void bits_decode_rows(uint32_t planes, uint32_t plane_words, uint64_t q_begin, uint64_t q_end, const uint32_t* win,
                      const uint32_t* counts, const uint32_t* bits, const uint64_t* out_off, uint32_t* out_ids,
                      uint8_t* out_hop, int32_t n_threads);

// traverse.cu: the k-hop batch on device buffers, on c->stream; counters land in c->h_counters once the
// caller has synchronised.  Uses c->status / scratch / tmp0..tmp3.
int32_t run_khop(CallCtx* c, dmg_graph* g, int mode, int k, int exclude, uint64_t nq, const uint32_t* d_q_off,
                 const uint32_t* d_q_src, uint64_t* d_out_off, uint32_t* d_out_ids, uint8_t* d_out_hop, uint64_t cap);

// NVTX ranges (tracing): with DMG_NVTX=1 in the environment every launch scope and every
// call is a named range on the host thread (nvtx3 is header-only and inert without a profiler attached).
bool nvtx_enabled();
void nvtx_push(const char* name);
void nvtx_pop();

// Launch helper: counts the launch, brackets it with events when profiling is on.
struct LaunchScope {
  CallCtx* c;
  cudaEvent_t a = nullptr, b = nullptr;
  int kid;
  bool nvtx = false;
  LaunchScope(CallCtx* ctx, int kernel_id, bool count = true, int launches = 1) : c(ctx), kid(kernel_id) {
    if (nvtx_enabled()) {
      nvtx_push(kKernelNames[kernel_id]);
      nvtx = true;
    }
    if (count) c->stats.kernel_launches += launches;
    if (c->eng->profile.load(std::memory_order_relaxed)) {
      a = c->get_event();
      b = c->get_event();
      cudaEventRecord(a, c->stream);
    }
  }
  ~LaunchScope() {
    if (a) {
      cudaEventRecord(b, c->stream);
      c->prof.push_back({kid, a, b});
    }
    if (nvtx) nvtx_pop();
  }
};

}  // namespace dmg

// ---------------------------------------------------------------------------------------------
// opaque handle bodies
// ---------------------------------------------------------------------------------------------
struct dmg_engine : dmg::Engine {};

struct dmg_graph {
  dmg_engine* eng = nullptr;
  std::atomic<int> refs{1};
  dmg_graph_info info{};
  // device-resident CSR (u32 offsets: E < 2^32), SURVEY.md 8 conventions
  uint32_t* d_out_ptr = nullptr;   // [V+1]
  uint32_t* d_out_idx = nullptr;   // [E]   row order = addDependency order
  uint32_t* d_in_ptr = nullptr;    // [V+1]
  uint32_t* d_in_idx = nullptr;    // [E]   row order = ascending source id
  uint32_t* d_und_ptr = nullptr;   // [V+1] resolve() rows: out row followed by in row (PG:359-378)
  uint32_t* d_und_idx = nullptr;   // [2E]
  uint32_t* d_part_off = nullptr;  // [P+1]
  uint32_t* d_entry = nullptr;     // [n_entry] insertion order
  uint32_t* d_entry_by_part = nullptr;   // [n_entry] grouped by project, insertion order inside each
  uint32_t* d_entry_bits = nullptr;  // [ceil(V/32)]
  uint8_t* d_vmeta = nullptr;        // [V] bits 0-6 host class-type code (0 = none), bit 7 entry point
  uint64_t* d_name_off = nullptr;  // [V+1] or null
  uint8_t* d_lower = nullptr;      // pre-lowered name arena (+32 B zero pad)
  uint8_t* d_names = nullptr;      // original name arena (exact identifier lookups are case sensitive)
  uint32_t* d_full_tbl = nullptr;  // open addressing: hash(identifier) -> vertex id, 0xFFFFFFFF = empty
  uint32_t* d_simple_tbl = nullptr;  // hash(lower-cased simple name) -> FIRST vertex of each project with that simple name
  uint32_t* d_simple_next = nullptr; // next vertex (ascending id) of the same project with the same simple name
  uint32_t tbl_mask = 0;
  // Per-project sliced-ELL copies of the three row sets the bit-parallel k-hop pulls from (see
  // khop_bits.cuh); built only when every project has <= 4096 vertices.  [0] = resolve rows (mode
  // BOTH), [1] = in rows (mode OUT pulls through dependents), [2] = out rows (mode IN).
  struct Ell {
    uint32_t* d_proj_slice = nullptr;   // [P+1] first slice of each project
    uint32_t* d_slice_off = nullptr;    // [S+1] start of each slice in d_data, in uint2 units
    uint32_t* d_vrow = nullptr;         // [S*32] (forward degree << 16) | vertex - project base; 0xFFFF = no row
    uint2* d_data = nullptr;            // slice = L/4 x 32 lanes x 4 u16 ids (lane-major inside a step); pad id = win_cap
    uint64_t entries = 0;               // padded size in ids
  } ell[3];
  bool has_ell = false;
  uint32_t* d_method_off = nullptr;  // [V+1] optional method catalog (names only)
  uint32_t* d_mname_off = nullptr;   // [M+1]
  uint8_t* d_mname_lower = nullptr;
  uint32_t* d_mexact_off = nullptr;  // [M+1] optional: method names as written (stack-trace frames match them exactly)
  uint8_t* d_mexact = nullptr;
  uint32_t n_methods = 0;
  std::vector<uint32_t> h_endpoint_methods;   // catalog indices of HTTP endpoint methods, ascending
  std::vector<uint32_t> h_part_off, h_entry;
  std::vector<uint32_t> h_entry_part_off;   // [P+1] slices of d_entry_by_part
  std::vector<uint32_t> h_method_off;       // host copy of the method catalog offsets (project slices of the endpoint list)
};

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
namespace dmg {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ uint32_t warp_sum_u32(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ uint32_t warp_incl_scan_u32(uint32_t v) {
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(FULL, v, o);
    if (lane >= (unsigned)o) v += t;
  }
  return v;
}
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Ordered output allocation across work items that finish out of order (decoupled look-back).
// Work item `q` (items are handed out by an atomic ticket, so every predecessor of a running
// item has itself started) publishes how many entries it produced and learns the sum over all
// items before it.  status[] must be zero before the launch.  One 64-bit word carries flag and
// value, so a relaxed load sees both or neither.  Call with all 32 lanes of ONE warp.
constexpr uint64_t LB_AGG = 1ull << 62, LB_PRE = 2ull << 62, LB_VAL = (1ull << 62) - 1;

__device__ __forceinline__ uint64_t lookback_exclusive(uint64_t* status, uint64_t q, uint64_t count) {
  const unsigned lane = threadIdx.x & 31;
  if (q == 0) {
    if (lane == 0) st_relaxed_u64(status, LB_PRE | count);
    return 0;
  }
  if (lane == 0) st_relaxed_u64(status + q, LB_AGG | count);
  uint64_t excl = 0;
  long long idx = (long long)q - 1;
  while (true) {
    const long long j = idx - (long long)lane;
    uint64_t v;
    v = (j >= 0) ? ld_relaxed_u64(status + j) : LB_PRE;
    while (__any_sync(FULL, (v >> 62) == 0)) {
      __nanosleep(100);   // predecessors are still working: do not burn issue slots spinning
      v = (j >= 0) ? ld_relaxed_u64(status + j) : LB_PRE;
    }
    const unsigned pm = __ballot_sync(FULL, (v >> 62) == 2);
    if (pm) {
      const int first = __ffs(pm) - 1;
      excl += warp_sum_u64(((int)lane <= first) ? (v & LB_VAL) : 0ull);
      break;
    }
    excl += warp_sum_u64(v & LB_VAL);
    idx -= 32;
  }
  if (lane == 0) st_relaxed_u64(status + q, LB_PRE | (excl + count));
  return excl;
}

__device__ __forceinline__ unsigned long long warp_incl_scan_u64(unsigned long long v) {
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned long long t = __shfl_up_sync(FULL, v, o);
    if (lane >= (unsigned)o) v += t;
  }
  return v;
}

// 64-bit exclusive scan over the block (row sizes can sum past 2^32 inside one tile: a batch may
// repeat a hub id, and k-hop rows of a multi-million-vertex project are each millions long);
// s_wsum has THREADS/32 + 1 words; all threads must call
template <int THREADS>
__device__ __forceinline__ unsigned long long block_excl_scan_u64(unsigned long long v, unsigned long long* s_wsum,
                                                                  unsigned long long* total) {
  constexpr int NW = THREADS / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned long long inc = warp_incl_scan_u64(v);
  if (lane == 31) s_wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = lane < NW ? s_wsum[lane] : 0;
    const unsigned long long wi = warp_incl_scan_u64(w);
    if (lane < NW) s_wsum[lane] = wi - w;
    if (lane == NW - 1) s_wsum[NW] = wi;
  }
  __syncthreads();
  const unsigned long long r = s_wsum[warp] + inc - v;
  *total = s_wsum[NW];
  __syncthreads();
  return r;
}

template <int THREADS>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_wsum, uint32_t* total) {
  // exclusive scan over the block; s_wsum has THREADS/32 + 1 words; all threads must call
  constexpr int NW = THREADS / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t inc = warp_incl_scan_u32(v);
  if (lane == 31) s_wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < NW ? s_wsum[lane] : 0;
    const uint32_t wi = warp_incl_scan_u32(w);
    if (lane < NW) s_wsum[lane] = wi - w;
    if (lane == NW - 1) s_wsum[NW] = wi;
  }
  __syncthreads();
  const uint32_t r = s_wsum[warp] + inc - v;
  *total = s_wsum[NW];
  __syncthreads();
  return r;
}


// Exclusive scan of the row sizes into out_off (u64), tile by tile with the look-back
// allocator; *carry (device) holds the running total across query chunks.
constexpr int SC_THREADS = 256, SC_ITEMS = 8;
static __global__ void __launch_bounds__(SC_THREADS) scan_counts_kernel(const uint32_t* __restrict__ counts,
                                                                 unsigned long long n, unsigned long long* out_off,
                                                                 unsigned long long* carry, uint64_t* status,
                                                                 DevCounters* ctr) {
  __shared__ unsigned long long s_t, s_base;
  __shared__ unsigned long long s_wsum[SC_THREADS / 32 + 1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr unsigned long long TILE = (unsigned long long)SC_THREADS * SC_ITEMS;
  const unsigned long long n_tiles = (n + TILE - 1) / TILE;
  const unsigned long long carry_in = *carry;
  while (true) {
    if (tid == 0) s_t = atomicAdd(&ctr->aux[1], 1ull);
    __syncthreads();
    const unsigned long long t = s_t;
    if (t >= n_tiles) break;
    const unsigned long long i0 = t * TILE + (unsigned long long)tid * SC_ITEMS;
    uint32_t x[SC_ITEMS];
    unsigned long long sum = 0;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; ++j) {
      x[j] = (i0 + j < n) ? counts[i0 + j] : 0;
      sum += x[j];
    }
    unsigned long long total;
    const unsigned long long off = block_excl_scan_u64<SC_THREADS>(sum, s_wsum, &total);
    if (warp == 0) {
      const unsigned long long excl = lookback_exclusive(status, t, total);
      if (lane == 0) s_base = excl;
    }
    __syncthreads();
    unsigned long long o = carry_in + s_base + off;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; ++j) {
      if (i0 + j < n) out_off[i0 + j] = o;
      o += x[j];
    }
    if (t == n_tiles - 1 && tid == SC_THREADS - 1) {
      // the last thread of the last tile holds the grand total
      out_off[n] = o;
      ctr->total = o;
      __threadfence();
      *carry = o;
    }
    __syncthreads();
  }
}


}  // namespace dmg
#endif
