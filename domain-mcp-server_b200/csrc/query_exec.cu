// Q1 of SURVEY.md 8a: batched execution of tokenised graph_query operations
// (GraphQueryService.execute, GQS:53-73).  Operations are grouped by kind, each group runs as ONE
// batch on the device primitives (adjacency rows, k-hop, name scans, method lookup), and the rows
// are put back in operation order.
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
  uint32_t *d_mo = nullptr, *d_no = nullptr;
  uint8_t* d_b = nullptr;
  DMG_CUDA(cudaMalloc((void**)&d_mo, ((size_t)V + 1) * 4));
  DMG_CUDA(cudaMalloc((void**)&d_no, ((size_t)M + 1) * 4));
  DMG_CUDA(cudaMalloc((void**)&d_b, (size_t)bytes + 16));
  DMG_CUDA(cudaMemcpy(d_mo, method_off, ((size_t)V + 1) * 4, cudaMemcpyHostToDevice));
  DMG_CUDA(cudaMemcpy(d_no, name_off, ((size_t)M + 1) * 4, cudaMemcpyHostToDevice));
  if (bytes) DMG_CUDA(cudaMemcpy(d_b, lower_bytes, bytes, cudaMemcpyHostToDevice));
  if (g->d_method_off) cudaFree(g->d_method_off);
  if (g->d_mname_off) cudaFree(g->d_mname_off);
  if (g->d_mname_lower) cudaFree(g->d_mname_lower);
  g->d_method_off = d_mo;
  g->d_mname_off = d_no;
  g->d_mname_lower = d_b;
  g->n_methods = M;
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

int32_t dmg_query_batch(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                        uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint64_t cap, uint64_t* needed) {
  if (!g || !out_off || !needed || (n_ops && !ops)) {
    set_error("dmg_query_batch: NULL argument");
    return DMG_EINVAL;
  }
  *needed = 0;
  out_off[0] = 0;
  if (!n_ops) return DMG_OK;
  // ---- group the operations ----------------------------------------------------------------
  // group results are overwritten in full by the primitives: no zero fill of gigabytes
  auto raw_u32 = [](uint64_t n) { return std::unique_ptr<uint32_t[]>(new uint32_t[std::max<uint64_t>(n, 1)]); };
  auto raw_u8 = [](uint64_t n) { return std::unique_ptr<uint8_t[]>(new uint8_t[std::max<uint64_t>(n, 1)]); };
  std::vector<uint64_t> adj_out, adj_in, names, methods, entries, endpoints;
  struct KhopKey {
    uint32_t mode, k, excl;
    bool operator<(const KhopKey& o) const {
      return mode != o.mode ? mode < o.mode : (k != o.k ? k < o.k : excl < o.excl);
    }
  };
  std::vector<std::pair<KhopKey, uint64_t>> khops;
  for (uint64_t i = 0; i < n_ops; ++i) {
    const dmg_query_op& op = ops[i];
    switch (op.opcode) {
      case DMG_OP_DEPENDENCIES: adj_out.push_back(i); break;
      case DMG_OP_DEPENDENTS: adj_in.push_back(i); break;
      case DMG_OP_OVERVIEW: adj_out.push_back(i); adj_in.push_back(i); break;
      case DMG_OP_KHOP:
        if (op.mode > DMG_BOTH || op.k < 1 || op.k > DMG_MAX_HOPS) {
          set_error("dmg_query_batch: op %llu: bad k-hop mode/k", (unsigned long long)i);
          return DMG_EINVAL;
        }
        khops.push_back({{op.mode, op.k, op.exclude_sources ? 1u : 0u}, i});
        break;
      case DMG_OP_ENTRYPOINTS: entries.push_back(i); break;
      case DMG_OP_ENDPOINTS: endpoints.push_back(i); break;
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
  std::vector<uint64_t> cnt(n_ops, 0);
  // a call without output buffers only sizes the result: the groups' rows are counted, not fetched
  const bool sizing_only = !out_ids || !out_aux || cap == 0;
  // per-group results, kept until the rows are laid out in operation order
  std::vector<uint64_t> o_off, i_off, k_off_all;
  std::unique_ptr<uint32_t[]> o_ids, i_ids;
  uint64_t nd = 0;
  int32_t st;
  auto adjacency = [&](int dir, const std::vector<uint64_t>& who, std::vector<uint64_t>& off, std::unique_ptr<uint32_t[]>& ids) {
    if (who.empty()) return (int32_t)DMG_OK;
    std::vector<uint32_t> v(who.size());
    for (size_t j = 0; j < who.size(); ++j) v[j] = ops[who[j]].vertex;
    off.assign(who.size() + 1, 0);
    int32_t s2 = dmg_adjacency_batch(g, dir, who.size(), v.data(), off.data(), nullptr, 0, &nd);
    if (s2 != DMG_OK && s2 != DMG_ECAP) return s2;
    if (nd && !sizing_only) {
      ids = raw_u32(nd);
      s2 = dmg_adjacency_batch(g, dir, who.size(), v.data(), off.data(), ids.get(), nd, &nd);
    } else {
      s2 = DMG_OK;
    }
    for (size_t j = 0; j < who.size(); ++j) cnt[who[j]] += off[j + 1] - off[j];
    return s2;
  };
  if ((st = adjacency(DMG_OUT, adj_out, o_off, o_ids)) != DMG_OK) return st;
  if ((st = adjacency(DMG_IN, adj_in, i_off, i_ids)) != DMG_OK) return st;
  // k-hop groups: one batch per (mode, k, exclude)
  std::sort(khops.begin(), khops.end());
  struct KGroup {
    size_t begin, end;
    std::vector<uint64_t> off;
    std::unique_ptr<uint32_t[]> ids;
    std::unique_ptr<uint8_t[]> hop;
  };
  std::vector<KGroup> kgroups;
  for (size_t a = 0; a < khops.size();) {
    size_t b = a;
    while (b < khops.size() && !(khops[a].first < khops[b].first) && !(khops[b].first < khops[a].first)) ++b;
    KGroup kg;
    kg.begin = a;
    kg.end = b;
    const size_t m = b - a;
    std::vector<uint32_t> qo(m + 1), qs(m);
    for (size_t j = 0; j < m; ++j) {
      qo[j] = (uint32_t)j;
      qs[j] = ops[khops[a + j].second].vertex;
    }
    qo[m] = (uint32_t)m;
    kg.off.assign(m + 1, 0);
    const KhopKey& key = khops[a].first;
    st = dmg_neighbors_batch(g, key.mode, key.k, key.excl, m, qo.data(), qs.data(), kg.off.data(), nullptr, nullptr, 0, &nd);
    if (st != DMG_OK && st != DMG_ECAP) return st;
    if (nd && !sizing_only) {
      kg.ids = raw_u32(nd);
      kg.hop = raw_u8(nd);
      st = dmg_neighbors_batch(g, key.mode, key.k, key.excl, m, qo.data(), qs.data(), kg.off.data(), kg.ids.get(),
                               kg.hop.get(), nd, &nd);
      if (st != DMG_OK) return st;
    }
    for (size_t j = 0; j < m; ++j) cnt[khops[a + j].second] = kg.off[j + 1] - kg.off[j];
    kgroups.push_back(std::move(kg));
    a = b;
  }
  // name resolution: simple-name equality first, containment for the misses (GQS:431-453)
  std::vector<uint32_t> name_id(names.size(), 0xFFFFFFFFu);
  std::vector<uint8_t> name_how(names.size(), 0);
  if (!names.empty()) {
    if (!g->d_name_off) {
      set_error("dmg_query_batch: DMG_OP_RESOLVE_NAME needs a graph with names");
      return DMG_EINVAL;
    }
    std::vector<size_t> todo(names.size());
    for (size_t j = 0; j < names.size(); ++j) todo[j] = j;
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
      st = dmg_search_batch(g, pass == 0 ? DMG_SEARCH_SIMPLE_EQ_FIRST : DMG_SEARCH_CONTAINS_FIRST, 0,
                            g->info.n_vertices, todo.size(), qo.data(), qb.data(), off.data(), ids.data(),
                            todo.size(), &nd);
      if (st != DMG_OK) return st;
      std::vector<size_t> miss;
      for (size_t j = 0; j < todo.size(); ++j) {
        if (off[j + 1] > off[j]) {
          name_id[todo[j]] = ids[off[j]];
          name_how[todo[j]] = pass == 0 ? 2 : 3;
        } else {
          miss.push_back(todo[j]);
        }
      }
      todo.swap(miss);
    }
    for (size_t j = 0; j < names.size(); ++j) cnt[names[j]] = name_id[j] != 0xFFFFFFFFu;
  }
  // method lookups
  std::vector<uint32_t> method_idx(methods.size(), 0xFFFFFFFFu);
  if (!methods.empty()) {
    CtxGuard cg(g->eng);
    CallCtx* c = cg.c;
    if (!c) return DMG_ECUDA;
    DMG_TRY(c->begin());
    const size_t m = methods.size();
    std::vector<uint32_t> v(m), tb(m), te(m);
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
    DMG_TRY(c->finish());
    for (size_t j = 0; j < m; ++j) cnt[methods[j]] = method_idx[j] != 0xFFFFFFFFu;
  }
  for (uint64_t i : entries) cnt[i] = g->h_entry.size();
  for (uint64_t i : endpoints) cnt[i] = g->h_endpoint_methods.size();

  // ---- lay the rows out in operation order ---------------------------------------------------
  uint64_t total = 0;
  for (uint64_t i = 0; i < n_ops; ++i) {
    out_off[i] = total;
    total += cnt[i];
  }
  out_off[n_ops] = total;
  *needed = total;
  if (total > cap || (total && (!out_ids || !out_aux))) {
    set_error("dmg_query_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    return DMG_ECAP;
  }
  std::vector<uint64_t> cur(out_off, out_off + n_ops);
  auto put = [&](uint64_t op, const uint32_t* ids, uint64_t n, uint8_t aux, const uint8_t* auxv) {
    if (!n) return;
    uint64_t& w = cur[op];
    std::memcpy(out_ids + w, ids, n * sizeof(uint32_t));
    if (auxv) std::memcpy(out_aux + w, auxv, n);
    else std::memset(out_aux + w, aux, n);
    w += n;
  };
  for (size_t j = 0; j < adj_out.size(); ++j) put(adj_out[j], o_ids.get() + o_off[j], o_off[j + 1] - o_off[j], 0, nullptr);
  for (size_t j = 0; j < adj_in.size(); ++j)
    put(adj_in[j], i_ids.get() + i_off[j], i_off[j + 1] - i_off[j], ops[adj_in[j]].opcode == DMG_OP_OVERVIEW ? 1 : 0,
        nullptr);
  for (const KGroup& kg : kgroups)
    for (size_t j = 0; j + kg.begin < kg.end; ++j)
      put(khops[kg.begin + j].second, kg.ids.get() + kg.off[j], kg.off[j + 1] - kg.off[j], 0, kg.hop.get() + kg.off[j]);
  for (size_t j = 0; j < names.size(); ++j)
    if (name_id[j] != 0xFFFFFFFFu) put(names[j], &name_id[j], 1, name_how[j], nullptr);
  for (size_t j = 0; j < methods.size(); ++j)
    if (method_idx[j] != 0xFFFFFFFFu) put(methods[j], &method_idx[j], 1, 0, nullptr);
  for (uint64_t i : entries) put(i, g->h_entry.data(), g->h_entry.size(), 0, nullptr);
  for (uint64_t i : endpoints) put(i, g->h_endpoint_methods.data(), g->h_endpoint_methods.size(), 0, nullptr);
  return DMG_OK;
}

}  // extern "C"
