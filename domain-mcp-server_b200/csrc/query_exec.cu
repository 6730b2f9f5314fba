// Q1 of SURVEY.md 8a: batched execution of tokenised graph_query operations
// (GraphQueryService.execute, GQS:53-73).  k-hop operations run as one device batch per (mode, k,
// exclude) and their rows stay on the device; names and methods are resolved by the lookup / scan
// primitives; adjacency rows, the entry list and the endpoint list are read where they already
// live.  The packed result (rows in operation order) is assembled on the device and copied once.
#include <algorithm>
#include <cstring>
#include <memory>
#include <vector>

#include "common.cuh"

namespace dmg {

// first method of the vertex whose lower-cased name equals the pattern (GQS:331-335, 367-371)
__global__ void method_lookup_kernel(const uint32_t* __restrict__ method_off, const uint32_t* __restrict__ mname_off,
                                     const uint8_t* __restrict__ mname, uint32_t V, uint64_t n,
                                     const uint32_t* __restrict__ vertex, const uint32_t* __restrict__ t_begin,
                                     const uint32_t* __restrict__ t_end, const uint8_t* __restrict__ text,
                                     uint32_t* out) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t res = 0xFFFFFFFFu;
  const uint32_t v = vertex[i];
  if (v < V) {
    const uint32_t pb = t_begin[i], pn = t_end[i] - pb;
    for (uint32_t m = method_off[v]; m < method_off[v + 1]; ++m) {
      const uint32_t b = mname_off[m], ln = mname_off[m + 1] - b;
      if (ln != pn) continue;
      uint32_t j = 0;
      while (j < pn && mname[b + j] == text[pb + j]) ++j;
      if (j == pn) {
        res = m;
        break;
      }
    }
  }
  out[i] = res;
}

// ---- device-side row assembly -------------------------------------------------------------------
// Every operation's row is at most two segments of existing device memory (a CSR row, the entry
// list, a row of a k-hop group's buffers, a single resolved id): qx_describe_kernel writes the
// segments and the row length, scan_counts_kernel turns lengths into out_off, qx_fill_kernel copies
// the segments into the packed result, which leaves the device in ONE copy.
constexpr uint32_t QX_IDS_ARE_VERTICES = 0x80000000u;   // flag in QxSeg::aux_const (its low byte is the aux value)
struct QxSeg {
  const uint32_t* ids;
  const uint8_t* aux;   // nullptr: every entry gets aux_const
  uint32_t len, aux_const;
};
struct QxRow {
  QxSeg s[2];
};
struct QxArgs {
  const dmg_query_op* ops;
  const unsigned long long* ref_a;
  const uint32_t *ref_b, *ref_g;
  const uint32_t* const* grp_ids;
  const uint8_t* const* grp_hop;
  const uint32_t *out_ptr, *out_idx, *in_ptr, *in_idx;
  uint32_t V;
  const uint32_t* entry;       // entry points grouped by project (insertion order inside each)
  const uint32_t* endpoints;   // endpoint methods, ascending catalog index
  QxRow* rows;
  uint32_t* cnt;
  unsigned long long n;
};

__global__ void __launch_bounds__(256) qx_describe_kernel(const QxArgs a) {
  const unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const dmg_query_op op = a.ops[i];
  QxRow r;
  r.s[0] = QxSeg{nullptr, nullptr, 0u, 0u};
  r.s[1] = QxSeg{nullptr, nullptr, 0u, 0u};
  const bool known = op.vertex < a.V;   // unknown vertices give empty rows
  switch (op.opcode) {
    case DMG_OP_DEPENDENCIES:
    case DMG_OP_OVERVIEW:
      if (known) {
        const uint32_t b = a.out_ptr[op.vertex];
        r.s[0] = QxSeg{a.out_idx + b, nullptr, a.out_ptr[op.vertex + 1] - b, QX_IDS_ARE_VERTICES};
      }
      if (op.opcode == DMG_OP_DEPENDENCIES) break;
      // overview: dependencies (aux 0) then dependents (aux 1)
    case DMG_OP_DEPENDENTS:
      if (known) {
        const uint32_t b = a.in_ptr[op.vertex];
        r.s[op.opcode == DMG_OP_OVERVIEW ? 1 : 0] =
            QxSeg{a.in_idx + b, nullptr, a.in_ptr[op.vertex + 1] - b,
                  (op.opcode == DMG_OP_OVERVIEW ? 1u : 0u) | QX_IDS_ARE_VERTICES};
      }
      break;
    case DMG_OP_KHOP: {
      const uint32_t grp = a.ref_g[i];
      const unsigned long long at = a.ref_a[i];
      const uint32_t len = a.grp_ids[grp] ? a.ref_b[i] : 0u;
      r.s[0] = QxSeg{a.grp_ids[grp] + at, a.grp_hop[grp] + at, len, QX_IDS_ARE_VERTICES};
      if (!a.grp_ids[grp]) r.s[0] = QxSeg{nullptr, nullptr, a.ref_b[i], 0u};   // sizing call: length only
      break;
    }
    // the op's project owns a contiguous slice of each list (host: ref_a = first entry, ref_b = length)
    case DMG_OP_ENTRYPOINTS: r.s[0] = QxSeg{a.entry + a.ref_a[i], nullptr, a.ref_b[i], QX_IDS_ARE_VERTICES}; break;
    case DMG_OP_ENDPOINTS: r.s[0] = QxSeg{a.endpoints + a.ref_a[i], nullptr, a.ref_b[i], 0u}; break;
    case DMG_OP_RESOLVE_NAME:
    case DMG_OP_METHOD:
      // the resolved id sits in the low word of the op's ref_a slot
      r.s[0] = QxSeg{reinterpret_cast<const uint32_t*>(a.ref_a + i), nullptr, a.ref_b[i],
                     op.opcode == DMG_OP_RESOLVE_NAME ? (a.ref_g[i] | QX_IDS_ARE_VERTICES) : 0u};
      break;
    default: break;
  }
  a.rows[i] = r;
  a.cnt[i] = r.s[0].len + r.s[1].len;
}

// Work is split by OUTPUT position (rows run from nothing to a whole project): a warp takes 2048
// positions, finds the row of the first one by bisection over out_off and walks forward.
__global__ void __launch_bounds__(256) qx_fill_kernel(const QxRow* __restrict__ rows,
                                                      const unsigned long long* __restrict__ out_off,
                                                      unsigned long long n_ops, uint32_t* out_ids, uint8_t* out_aux,
                                                      unsigned long long total, const uint8_t* __restrict__ vmeta,
                                                      uint32_t V, uint8_t* out_meta) {
  constexpr unsigned long long CHUNK = 2048;
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
  for (unsigned long long P = warp * CHUNK; P < total; P += nwarps * CHUNK) {
    const unsigned long long Pend = min(total, P + CHUNK);
    unsigned long long q = 0, hi = n_ops;   // first row whose end lies beyond P (empty rows are skipped)
    while (q < hi) {
      const unsigned long long mid = (q + hi) >> 1;
      if (out_off[mid + 1] <= P) q = mid + 1; else hi = mid;
    }
    unsigned long long row_begin = out_off[q], row_end = out_off[q + 1];
    QxRow r = rows[q];
    for (unsigned long long p = P + lane; p < Pend; p += 32) {
      while (p >= row_end) {
        ++q;
        row_begin = row_end;
        row_end = out_off[q + 1];
        if (p < row_end) r = rows[q];
      }
      uint32_t j = (uint32_t)(p - row_begin);
      const int sg = j < r.s[0].len ? 0 : 1;
      if (sg) j -= r.s[0].len;
      const uint32_t id = r.s[sg].ids[j];
      out_ids[p] = id;
      out_aux[p] = r.s[sg].aux ? r.s[sg].aux[j] : (uint8_t)r.s[sg].aux_const;
      // decoration (GQS:122-130, 308-316): class-type code + entry flag of a vertex id; 0 for catalog indices
      if (out_meta) out_meta[p] = ((r.s[sg].aux_const & QX_IDS_ARE_VERTICES) && id < V) ? vmeta[id] : (uint8_t)0;
    }
  }
}

// device allocation that frees itself on every exit path of the executor; with an engine set, the
// block comes from and goes back to the engine's pool instead of cudaMalloc / cudaFree
struct DevMem {
  void* p = nullptr;
  size_t got = 0;
  Engine* pool = nullptr;
  ~DevMem() { reset(); }
  void reset() {
    if (p) {
      if (pool) pool->pool_give(p, got);
      else cudaFree(p);
    }
    p = nullptr;
  }
  int32_t alloc(size_t bytes) {
    reset();
    if (pool) {
      p = pool->pool_take(bytes, &got);
      if (!p) {
        set_error("dmg_query_batch: device out of memory (%zu bytes)", bytes);
        return DMG_ENOMEM;
      }
      return DMG_OK;
    }
    DMG_CUDA(cudaMalloc(&p, std::max<size_t>(bytes, 256)));
    return DMG_OK;
  }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace dmg

using namespace dmg;

extern "C" {

int32_t dmg_graph_set_methods(dmg_graph* g, const uint32_t* method_off, const uint32_t* name_off,
                              const uint8_t* lower_bytes) {
  if (!g || !method_off || !name_off) {
    set_error("dmg_graph_set_methods: NULL argument");
    return DMG_EINVAL;
  }
  const uint32_t V = g->info.n_vertices;
  for (uint32_t v = 0; v < V; ++v)
    if (method_off[v + 1] < method_off[v]) {
      set_error("dmg_graph_set_methods: method_off must be non-decreasing");
      return DMG_EINVAL;
    }
  const uint32_t M = method_off[V];
  const uint32_t bytes = name_off[M];
  if (bytes && !lower_bytes) {
    set_error("dmg_graph_set_methods: lower_bytes is NULL");
    return DMG_EINVAL;
  }
  DMG_CUDA(cudaSetDevice(g->eng->device));
  DevMem mo, no, lb;   // freed on every failure path; handed to the graph only when all three are filled
  DMG_TRY(mo.alloc(((size_t)V + 1) * 4));
  DMG_TRY(no.alloc(((size_t)M + 1) * 4));
  DMG_TRY(lb.alloc((size_t)bytes + 16));
  DMG_CUDA(cudaMemcpy(mo.p, method_off, ((size_t)V + 1) * 4, cudaMemcpyHostToDevice));
  DMG_CUDA(cudaMemcpy(no.p, name_off, ((size_t)M + 1) * 4, cudaMemcpyHostToDevice));
  if (bytes) DMG_CUDA(cudaMemcpy(lb.p, lower_bytes, bytes, cudaMemcpyHostToDevice));
  if (g->d_method_off) cudaFree(g->d_method_off);
  if (g->d_mname_off) cudaFree(g->d_mname_off);
  if (g->d_mname_lower) cudaFree(g->d_mname_lower);
  g->d_method_off = mo.as<uint32_t>();
  g->d_mname_off = no.as<uint32_t>();
  g->d_mname_lower = lb.as<uint8_t>();
  mo.p = no.p = lb.p = nullptr;   // ownership moved to the graph
  g->n_methods = M;
  try {
    g->h_method_off.assign(method_off, method_off + (size_t)V + 1);
  } catch (const std::exception&) {
    set_error("dmg_graph_set_methods: host out of memory");
    return DMG_ENOMEM;
  }
  return DMG_OK;
}

int32_t dmg_graph_set_method_flags(dmg_graph* g, const uint8_t* flags) {
  if (!g || !g->d_method_off || (g->n_methods && !flags)) {
    set_error("dmg_graph_set_method_flags: needs dmg_graph_set_methods first and a flag per method");
    return DMG_EINVAL;
  }
  g->h_endpoint_methods.clear();
  for (uint32_t m = 0; m < g->n_methods; ++m)
    if (flags[m] & 1u) g->h_endpoint_methods.push_back(m);   // MethodInfo.isHttpEndpoint(), PG:75-77
  return DMG_OK;
}

static int32_t query_batch_impl(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                                uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint8_t* out_meta, uint64_t cap,
                                uint64_t* needed);
int32_t dmg_query_batch(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                        uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint64_t cap, uint64_t* needed) {
  return query_batch_impl(g, n_ops, ops, text_bytes, out_off, out_ids, out_aux, nullptr, cap, needed);
}
int32_t dmg_query_batch_meta(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                             uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint8_t* out_meta, uint64_t cap,
                             uint64_t* needed) {
  return query_batch_impl(g, n_ops, ops, text_bytes, out_off, out_ids, out_aux, out_meta, cap, needed);
}
int32_t dmg_graph_set_class_types(dmg_graph* g, const uint8_t* codes) {
  if (!g || (g->info.n_vertices && !codes)) {
    set_error("dmg_graph_set_class_types: NULL argument");
    return DMG_EINVAL;
  }
  const uint32_t V = g->info.n_vertices;
  std::vector<uint8_t> meta(std::max<uint32_t>(V, 1), 0);
  for (uint32_t v = 0; v < V; ++v) {
    if (codes[v] > 127) {
      set_error("dmg_graph_set_class_types: code %u of vertex %u exceeds 127", codes[v], v);
      return DMG_EINVAL;
    }
    meta[v] = codes[v];
  }
  for (uint32_t v : g->h_entry) meta[v] |= 0x80u;   // isEntryPoint (PG:336-338)
  DMG_CUDA(cudaSetDevice(g->eng->device));
  DMG_CUDA(cudaMemcpy(g->d_vmeta, meta.data(), meta.size(), cudaMemcpyHostToDevice));
  return DMG_OK;
}
static int32_t query_batch_impl(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                                uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint8_t* out_meta, uint64_t cap,
                                uint64_t* needed) {
  if (!g || !out_off || !needed || (n_ops && !ops)) {
    set_error("dmg_query_batch: NULL argument");
    return DMG_EINVAL;
  }
  *needed = 0;
  out_off[0] = 0;
  if (!n_ops) return DMG_OK;
  if (n_ops >= (1ull << 32)) {
    set_error("dmg_query_batch: too many operations");
    return DMG_EINVAL;
  }
  // a call without output buffers only sizes the result: k-hop groups are counted, not fetched
  const bool sizing_only = !out_ids || !out_aux || cap == 0;
  // ---- group the operations ----------------------------------------------------------------
  std::vector<uint64_t> names, methods;
  struct KhopKey {
    uint32_t mode, k, excl;
    bool operator<(const KhopKey& o) const {
      return mode != o.mode ? mode < o.mode : (k != o.k ? k < o.k : excl < o.excl);
    }
  };
  std::vector<std::pair<KhopKey, uint64_t>> khops;
  bool want_endpoints = false;
  const uint32_t n_parts = g->info.n_parts;
  for (uint64_t i = 0; i < n_ops; ++i) {
    const dmg_query_op& op = ops[i];
    // The reference answers graph_query from ONE project's graph (GQS:56-62): the three operations
    // that range over a whole graph are scoped to the op's project of this handle.
    if ((op.opcode == DMG_OP_ENTRYPOINTS || op.opcode == DMG_OP_ENDPOINTS || op.opcode == DMG_OP_RESOLVE_NAME) &&
        op.project >= n_parts) {
      set_error("dmg_query_batch: op %llu: project %u out of range (graph holds %u)", (unsigned long long)i,
                op.project, n_parts);
      return DMG_EINVAL;
    }
    switch (op.opcode) {
      case DMG_OP_DEPENDENCIES:
      case DMG_OP_DEPENDENTS:
      case DMG_OP_OVERVIEW:
      case DMG_OP_ENTRYPOINTS: break;
      case DMG_OP_ENDPOINTS: want_endpoints = true; break;
      case DMG_OP_KHOP:
        if (op.mode > DMG_BOTH || op.k < 1 || op.k > DMG_MAX_HOPS) {
          set_error("dmg_query_batch: op %llu: bad k-hop mode/k", (unsigned long long)i);
          return DMG_EINVAL;
        }
        khops.push_back({{op.mode, op.k, op.exclude_sources ? 1u : 0u}, i});
        break;
      case DMG_OP_RESOLVE_NAME:
      case DMG_OP_METHOD:
        if (op.text_end < op.text_begin || (op.text_end > op.text_begin && !text_bytes)) {
          set_error("dmg_query_batch: op %llu: bad text range", (unsigned long long)i);
          return DMG_EINVAL;
        }
        (op.opcode == DMG_OP_RESOLVE_NAME ? names : methods).push_back(i);
        break;
      default:
        set_error("dmg_query_batch: op %llu: unknown opcode %u", (unsigned long long)i, op.opcode);
        return DMG_EINVAL;
    }
  }
  if (!methods.empty() && !g->d_method_off) {
    set_error("dmg_query_batch: DMG_OP_METHOD needs dmg_graph_set_methods");
    return DMG_EINVAL;
  }
  DMG_CUDA(cudaSetDevice(g->eng->device));
  // what the host learns about an op before the rows are described on the device:
  //   k-hop: ref_a = start of its row in the group's buffers, ref_b = row length, ref_g = group
  //   name / method: ref_a = the resolved id, ref_b = 0 or 1, ref_g = aux value
  std::vector<unsigned long long> ref_a(n_ops, 0);
  std::vector<uint32_t> ref_b(n_ops, 0), ref_g(n_ops, 0);
  uint64_t nd = 0;
  int32_t st;

  // ---- k-hop groups: one device batch per (mode, k, exclude); the rows stay on the device -----
  std::sort(khops.begin(), khops.end());
  struct KGroup {
    DevMem q, off, ids, hop;
  };
  std::vector<std::unique_ptr<KGroup>> kgroups;
  for (size_t a = 0; a < khops.size();) {
    size_t b = a;
    while (b < khops.size() && !(khops[a].first < khops[b].first) && !(khops[b].first < khops[a].first)) ++b;
    const size_t m = b - a;
    std::unique_ptr<KGroup> kg(new KGroup);
    kg->q.pool = kg->off.pool = kg->ids.pool = kg->hop.pool = g->eng;
    std::vector<uint32_t> q(2 * m + 1);   // q_off = 0..m, then q_src
    for (size_t j = 0; j <= m; ++j) q[j] = (uint32_t)j;
    for (size_t j = 0; j < m; ++j) q[m + 1 + j] = ops[khops[a + j].second].vertex;
    DMG_TRY(kg->q.alloc(q.size() * 4));
    DMG_TRY(kg->off.alloc((m + 1) * 8));
    DMG_CUDA(cudaMemcpy(kg->q.p, q.data(), q.size() * 4, cudaMemcpyHostToDevice));
    const KhopKey& key = khops[a].first;
    const uint32_t* d_qo = kg->q.as<uint32_t>();
    // A group can never hold more than the caller's capacity (or the batch is too large anyway)
    // nor more than a project per query: with buffers of that size ONE traversal both sizes and
    // fetches the group.  Only a guess cut short by free memory needs a second, exactly sized run.
    uint64_t bound = sizing_only ? 0 : cap;
    const uint64_t maxp = g->info.max_part_vertices;
    if (maxp && m <= UINT64_MAX / maxp) bound = std::min<uint64_t>(bound, m * maxp);
    uint64_t guess = bound;
    if (guess) {
      size_t free_b = 0, total_b = 0;
      DMG_CUDA(cudaMemGetInfo(&free_b, &total_b));
      free_b += g->eng->pool_bytes;   // kept blocks are there to be reused
      guess = std::min<uint64_t>(guess, free_b / 20);   // ids + hops here, the traversal's temporaries, and headroom
      guess = std::min<uint64_t>(guess, 1ull << 30);     // a caller with an "unbounded" capacity does not get 40 GB of guess
    }
    for (int attempt = 0; attempt < 2; ++attempt) {
      if (guess) {
        DMG_TRY(kg->ids.alloc(guess * 4));
        DMG_TRY(kg->hop.alloc(guess));
      }
      st = dmg_neighbors_batch_dev(g, key.mode, key.k, key.excl, m, d_qo, d_qo + m + 1, kg->off.as<uint64_t>(),
                                   kg->ids.as<uint32_t>(), kg->hop.as<uint8_t>(), guess, &nd);
      if (st == DMG_OK) break;
      if (st != DMG_ECAP) return st;
      kg->ids.reset();   // rows did not fit: the offsets are valid, the buffers are not
      kg->hop.reset();
      if (attempt == 1 || nd > bound) break;   // sizing call, or more than the caller can take: count only
      guess = nd;
    }
    std::vector<uint64_t> off(m + 1);
    DMG_CUDA(cudaMemcpy(off.data(), kg->off.p, (m + 1) * 8, cudaMemcpyDeviceToHost));
    for (size_t j = 0; j < m; ++j) {
      const uint64_t op = khops[a + j].second;
      ref_a[op] = off[j];
      ref_b[op] = (uint32_t)(off[j + 1] - off[j]);
      ref_g[op] = (uint32_t)kgroups.size();
    }
    kgroups.push_back(std::move(kg));
    a = b;
  }

  // ---- name resolution: simple-name equality first, containment for the misses (GQS:431-453) ---
  if (!names.empty()) {
    if (!g->d_name_off) {
      set_error("dmg_query_batch: DMG_OP_RESOLVE_NAME needs a graph with names");
      return DMG_EINVAL;
    }
    // one scan per project that has name ops (a scan covers one vertex range)
    std::stable_sort(names.begin(), names.end(),
                     [&](uint64_t x, uint64_t y) { return ops[x].project < ops[y].project; });
    for (size_t n0 = 0; n0 < names.size();) {
      size_t n1 = n0;
      const uint32_t proj = ops[names[n0]].project;
      while (n1 < names.size() && ops[names[n1]].project == proj) ++n1;
      const uint32_t v_lo = g->h_part_off[proj], v_hi = g->h_part_off[proj + 1];
      std::vector<size_t> todo(n1 - n0);
      for (size_t j = 0; j < todo.size(); ++j) todo[j] = n0 + j;
      for (int pass = 0; pass < 2 && !todo.empty(); ++pass) {
        std::vector<uint32_t> qo(todo.size() + 1, 0);
        std::vector<uint8_t> qb;
        for (size_t j = 0; j < todo.size(); ++j) {
          const dmg_query_op& op = ops[names[todo[j]]];
          qb.insert(qb.end(), text_bytes + op.text_begin, text_bytes + op.text_end);
          qo[j + 1] = (uint32_t)qb.size();
        }
        qb.push_back(0);
        std::vector<uint64_t> off(todo.size() + 1, 0);
        std::vector<uint32_t> ids(todo.size() + 1, 0);
        st = dmg_search_batch(g, pass == 0 ? DMG_SEARCH_SIMPLE_EQ_FIRST : DMG_SEARCH_CONTAINS_FIRST, v_lo, v_hi,
                              todo.size(), qo.data(), qb.data(), off.data(), ids.data(), todo.size(), &nd);
        if (st != DMG_OK) return st;
        std::vector<size_t> miss;
        for (size_t j = 0; j < todo.size(); ++j) {
          if (off[j + 1] > off[j]) {
            const uint64_t op = names[todo[j]];
            ref_a[op] = ids[off[j]];
            ref_b[op] = 1;
            ref_g[op] = pass == 0 ? 2 : 3;
          } else {
            miss.push_back(todo[j]);
          }
        }
        todo.swap(miss);
      }
      n0 = n1;
    }
  }

  // ---- entry points / endpoints: the slice of the op's project ---------------------------------
  for (uint64_t i = 0; i < n_ops; ++i) {
    const dmg_query_op& op = ops[i];
    if (op.opcode == DMG_OP_ENTRYPOINTS) {
      ref_a[i] = g->h_entry_part_off[op.project];
      ref_b[i] = g->h_entry_part_off[op.project + 1] - g->h_entry_part_off[op.project];
    } else if (op.opcode == DMG_OP_ENDPOINTS && !g->h_method_off.empty()) {
      // catalog indices ascend with the vertex id, so a project's endpoints are one run of the list
      const uint32_t m_lo = g->h_method_off[g->h_part_off[op.project]], m_hi = g->h_method_off[g->h_part_off[op.project + 1]];
      const auto& ep = g->h_endpoint_methods;
      const size_t a0 = std::lower_bound(ep.begin(), ep.end(), m_lo) - ep.begin();
      const size_t a1 = std::lower_bound(ep.begin(), ep.end(), m_hi) - ep.begin();
      ref_a[i] = a0;
      ref_b[i] = (uint32_t)(a1 - a0);
    }
  }

  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  // ---- method lookups ------------------------------------------------------------------------
  if (!methods.empty()) {
    const size_t m = methods.size();
    std::vector<uint32_t> v(m), tb(m), te(m), method_idx(m, 0xFFFFFFFFu);
    std::vector<uint8_t> txt;
    for (size_t j = 0; j < m; ++j) {
      const dmg_query_op& op = ops[methods[j]];
      v[j] = op.vertex;
      tb[j] = (uint32_t)txt.size();
      txt.insert(txt.end(), text_bytes + op.text_begin, text_bytes + op.text_end);
      te[j] = (uint32_t)txt.size();
    }
    txt.push_back(0);
    DMG_TRY(c->tmp0.reserve(m * 16 + txt.size() + 64));
    uint32_t* d_v = c->tmp0.as<uint32_t>();
    uint32_t *d_tb = d_v + m, *d_te = d_tb + m, *d_out = d_te + m;
    uint8_t* d_txt = reinterpret_cast<uint8_t*>(d_out + m);
    DMG_CUDA(cudaMemcpyAsync(d_v, v.data(), m * 4, cudaMemcpyHostToDevice, c->stream));
    DMG_CUDA(cudaMemcpyAsync(d_tb, tb.data(), m * 4, cudaMemcpyHostToDevice, c->stream));
    DMG_CUDA(cudaMemcpyAsync(d_te, te.data(), m * 4, cudaMemcpyHostToDevice, c->stream));
    DMG_CUDA(cudaMemcpyAsync(d_txt, txt.data(), txt.size(), cudaMemcpyHostToDevice, c->stream));
    {
      LaunchScope ls(c, K_MISC);
      method_lookup_kernel<<<(unsigned)((m + 255) / 256), 256, 0, c->stream>>>(
          g->d_method_off, g->d_mname_off, g->d_mname_lower, g->info.n_vertices, m, d_v, d_tb, d_te, d_txt, d_out);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(method_idx.data(), d_out, m * 4, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    for (size_t j = 0; j < m; ++j)
      if (method_idx[j] != 0xFFFFFFFFu) {
        ref_a[methods[j]] = method_idx[j];
        ref_b[methods[j]] = 1;
      }
  }

  // ---- describe every row on the device, scan the sizes, fill the packed result -----------------
  const size_t n_grp = kgroups.size();
  std::vector<const uint32_t*> grp_ids(std::max<size_t>(n_grp, 1), nullptr);
  std::vector<const uint8_t*> grp_hop(std::max<size_t>(n_grp, 1), nullptr);
  for (size_t j = 0; j < n_grp; ++j) {
    grp_ids[j] = kgroups[j]->ids.as<uint32_t>();
    grp_hop[j] = kgroups[j]->hop.as<uint8_t>();
  }
  const size_t n_endp = want_endpoints ? g->h_endpoint_methods.size() : 0;
  // one upload buffer: ops | ref_a | ref_b | ref_g | group pointers | endpoint list
  const size_t o_ops = 0, o_ra = o_ops + n_ops * sizeof(dmg_query_op), o_rb = o_ra + n_ops * 8, o_rg = o_rb + n_ops * 4,
               o_gi = (o_rg + n_ops * 4 + 15) & ~size_t(15), o_gh = o_gi + grp_ids.size() * 8,
               o_ep = o_gh + grp_hop.size() * 8, up_bytes = o_ep + n_endp * 4;
  DMG_TRY(c->q_src.reserve(up_bytes));
  uint8_t* up = c->q_src.as<uint8_t>();
  DMG_CUDA(cudaMemcpyAsync(up + o_ops, ops, n_ops * sizeof(dmg_query_op), cudaMemcpyHostToDevice, c->stream));
  DMG_CUDA(cudaMemcpyAsync(up + o_ra, ref_a.data(), n_ops * 8, cudaMemcpyHostToDevice, c->stream));
  DMG_CUDA(cudaMemcpyAsync(up + o_rb, ref_b.data(), n_ops * 4, cudaMemcpyHostToDevice, c->stream));
  DMG_CUDA(cudaMemcpyAsync(up + o_rg, ref_g.data(), n_ops * 4, cudaMemcpyHostToDevice, c->stream));
  DMG_CUDA(cudaMemcpyAsync(up + o_gi, grp_ids.data(), grp_ids.size() * 8, cudaMemcpyHostToDevice, c->stream));
  DMG_CUDA(cudaMemcpyAsync(up + o_gh, grp_hop.data(), grp_hop.size() * 8, cudaMemcpyHostToDevice, c->stream));
  if (n_endp)
    DMG_CUDA(cudaMemcpyAsync(up + o_ep, g->h_endpoint_methods.data(), n_endp * 4, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += up_bytes;
  const uint64_t scan_tiles = (n_ops + SC_THREADS * SC_ITEMS - 1) / (SC_THREADS * SC_ITEMS);
  DMG_TRY(c->tmp1.reserve(n_ops * sizeof(QxRow)));
  DMG_TRY(c->tmp2.reserve(n_ops * 4));
  DMG_TRY(c->out_off.reserve((n_ops + 1) * 8));
  DMG_TRY(c->status.reserve(scan_tiles * 8));
  DMG_CUDA(cudaMemsetAsync(c->status.p, 0, scan_tiles * 8, c->stream));
  DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
  QxArgs qa;
  qa.ops = reinterpret_cast<const dmg_query_op*>(up + o_ops);
  qa.ref_a = reinterpret_cast<const unsigned long long*>(up + o_ra);
  qa.ref_b = reinterpret_cast<const uint32_t*>(up + o_rb);
  qa.ref_g = reinterpret_cast<const uint32_t*>(up + o_rg);
  qa.grp_ids = reinterpret_cast<const uint32_t* const*>(up + o_gi);
  qa.grp_hop = reinterpret_cast<const uint8_t* const*>(up + o_gh);
  qa.out_ptr = g->d_out_ptr;
  qa.out_idx = g->d_out_idx;
  qa.in_ptr = g->d_in_ptr;
  qa.in_idx = g->d_in_idx;
  qa.V = g->info.n_vertices;
  qa.entry = g->d_entry_by_part;
  qa.endpoints = reinterpret_cast<const uint32_t*>(up + o_ep);
  qa.rows = c->tmp1.as<QxRow>();
  qa.cnt = c->tmp2.as<uint32_t>();
  qa.n = n_ops;
  {
    LaunchScope ls(c, K_QUERY_ROWS);
    qx_describe_kernel<<<(unsigned)((n_ops + 255) / 256), 256, 0, c->stream>>>(qa);
  }
  {
    LaunchScope ls(c, K_SCAN_COUNTS);
    scan_counts_kernel<<<(unsigned)std::min<uint64_t>(scan_tiles, (uint64_t)g->eng->sm_count * 4), SC_THREADS, 0, c->stream>>>(
        qa.cnt, n_ops, c->out_off.as<unsigned long long>(), &c->d_counters->aux[2], c->status.as<uint64_t>(),
        c->d_counters);
  }
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaMemcpyAsync(out_off, c->out_off.p, (n_ops + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  const uint64_t total = c->h_counters->total;
  c->stats.d2h_bytes += (n_ops + 1) * 8 + sizeof(DevCounters);
  c->stats.results = total;
  *needed = total;
  if (total > cap || (total && sizing_only)) {
    DMG_TRY(c->finish());
    set_error("dmg_query_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    return DMG_ECAP;
  }
  if (total) {
    DMG_TRY(c->out_ids.reserve(total * 4));
    DMG_TRY(c->out_hop.reserve(total));
    if (out_meta) DMG_TRY(c->tmp3.reserve(total));
    {
      LaunchScope ls(c, K_QUERY_ROWS);
      qx_fill_kernel<<<(unsigned)g->eng->sm_count * 8, 256, 0, c->stream>>>(
          qa.rows, c->out_off.as<unsigned long long>(), n_ops, c->out_ids.as<uint32_t>(), c->out_hop.as<uint8_t>(), total,
          g->d_vmeta, g->info.n_vertices, out_meta ? c->tmp3.as<uint8_t>() : nullptr);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(out_ids, c->out_ids.p, total * 4, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaMemcpyAsync(out_aux, c->out_hop.p, total, cudaMemcpyDeviceToHost, c->stream));
    if (out_meta) DMG_CUDA(cudaMemcpyAsync(out_meta, c->tmp3.p, total, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += total * (out_meta ? 6 : 5);
  }
  return c->finish();
}

}  // extern "C"
