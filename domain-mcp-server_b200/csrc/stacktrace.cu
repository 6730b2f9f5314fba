// get_stack_trace_context, graph half, as ONE call (CCS:1392-1446 + 1550-1578):
// frame class names -> vertex ids (hash probe of the identifier table), frame status, the project of
// the first found frame, the found frames of that project as the source set, k-hop neighbour rows --
// all on the device, no id round trip through the host between lookup and traversal.
#include <algorithm>

#include "common.cuh"

namespace dmg {

struct StArgs {
  const unsigned long long* name_off;
  const uint8_t* names;        // identifiers as written (nodes.containsKey is case sensitive, PG:464)
  const uint32_t* full_tbl;
  uint32_t mask, V, n_parts;
  const uint32_t* part_off;
  const uint32_t* method_off;  // catalog (optional): methods of vertex v = [method_off[v], method_off[v+1])
  const uint32_t* mexact_off;
  const uint8_t* mexact;       // method names as written
  unsigned long long n_frames, n_traces;
  const uint32_t* t_off;       // [n_traces + 1] frames of a trace
  const uint32_t* cn_off;      // [n_frames + 1] class-name text of a frame
  const uint8_t* cn;
  const uint32_t* mn_off;      // method-name text of a frame (nullptr: no method check)
  const uint8_t* mn;
  uint8_t* status;             // [n_frames]
  uint32_t* vertex;            // [n_frames]
  uint32_t* method;            // [n_frames] catalog index, 0xFFFFFFFF = none
  uint32_t* q_src;             // [n_frames] the k-hop source of the frame, 0xFFFFFFFF = none
};

// smallest vertex of [lo, hi) whose identifier equals the text (the same FQCN may exist in two projects)
__device__ __forceinline__ uint32_t st_find(const StArgs& a, const uint8_t* q, uint32_t qn, uint32_t lo, uint32_t hi) {
  uint32_t best = 0xFFFFFFFFu;
  for (uint32_t h = name_hash(q, qn) & a.mask;; h = (h + 1) & a.mask) {
    const uint32_t v = a.full_tbl[h];
    if (v == 0xFFFFFFFFu) break;
    if (v < lo || v >= hi || v >= best) continue;
    const unsigned long long b = a.name_off[v];
    if (a.name_off[v + 1] - b != qn) continue;
    uint32_t j = 0;
    while (j < qn && a.names[b + j] == q[j]) ++j;
    if (j == qn) best = v;
  }
  return best;
}

// one thread per frame: class lookup over every project, then the method by name inside the class
__global__ void __launch_bounds__(256) st_frames_kernel(const StArgs a) {
  const unsigned long long f = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (f >= a.n_frames) return;
  const uint32_t v = st_find(a, a.cn + a.cn_off[f], a.cn_off[f + 1] - a.cn_off[f], 0u, a.V);
  uint32_t st = DMG_FRAME_MISSING, m_at = 0xFFFFFFFFu;
  if (v != 0xFFFFFFFFu) {
    st = DMG_FRAME_FOUND;
    if (a.mn_off && a.mexact_off) {   // findByClassNameAndMethodName: SQL equality on both (CCS:1415-1417)
      st = DMG_FRAME_CLASS_MISSING;
      const uint8_t* q = a.mn + a.mn_off[f];
      const uint32_t qn = a.mn_off[f + 1] - a.mn_off[f];
      for (uint32_t m = a.method_off[v]; m < a.method_off[v + 1]; ++m) {
        const uint32_t b = a.mexact_off[m];
        if (a.mexact_off[m + 1] - b != qn) continue;
        uint32_t j = 0;
        while (j < qn && a.mexact[b + j] == q[j]) ++j;
        if (j == qn) {
          st = DMG_FRAME_FOUND;
          m_at = m;
          break;
        }
      }
    }
  }
  a.status[f] = (uint8_t)st;
  a.vertex[f] = v;
  a.method[f] = m_at;
}

// one thread per trace: the project of the first found frame (CCS:1426-1434) is the graph of the trace;
// a found frame of another project takes part by NAME (matchedClasses holds names, CCS:1563-1568), so its
// class is looked up again inside that project.
__global__ void __launch_bounds__(128) st_sources_kernel(const StArgs a) {
  const unsigned long long t = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (t >= a.n_traces) return;
  uint32_t lo = 0, hi = 0;
  bool have = false;
  for (uint32_t f = a.t_off[t]; f < a.t_off[t + 1]; ++f) {
    uint32_t src = 0xFFFFFFFFu;
    if (a.status[f] == DMG_FRAME_FOUND) {
      const uint32_t v = a.vertex[f];
      if (!have) {
        uint32_t l = 0, r = a.n_parts;
        while (r - l > 1) {
          const uint32_t m = (l + r) >> 1;
          if (a.part_off[m] <= v) l = m; else r = m;
        }
        lo = a.part_off[l];
        hi = a.part_off[l + 1];
        have = true;
      }
      src = (v >= lo && v < hi) ? v : st_find(a, a.cn + a.cn_off[f], a.cn_off[f + 1] - a.cn_off[f], lo, hi);
    }
    a.q_src[f] = src;
  }
}

}  // namespace dmg

using namespace dmg;

extern "C" {

int32_t dmg_graph_set_method_names(dmg_graph* g, const uint32_t* name_off, const uint8_t* bytes) {
  if (!g || !name_off || !g->d_method_off) {
    set_error("dmg_graph_set_method_names: needs dmg_graph_set_methods first and the name offsets");
    return DMG_EINVAL;
  }
  const uint32_t M = g->n_methods, nb = name_off[M];
  if (nb && !bytes) {
    set_error("dmg_graph_set_method_names: bytes is NULL");
    return DMG_EINVAL;
  }
  DMG_CUDA(cudaSetDevice(g->eng->device));
  uint32_t* d_off = nullptr;
  uint8_t* d_b = nullptr;
  DMG_CUDA(cudaMalloc((void**)&d_off, ((size_t)M + 1) * 4));
  if (cudaMalloc((void**)&d_b, (size_t)nb + 16) != cudaSuccess) {
    cudaFree(d_off);
    set_error("dmg_graph_set_method_names: out of device memory");
    return DMG_ENOMEM;
  }
  cudaMemcpy(d_off, name_off, ((size_t)M + 1) * 4, cudaMemcpyHostToDevice);
  if (nb) cudaMemcpy(d_b, bytes, nb, cudaMemcpyHostToDevice);
  if (g->d_mexact_off) cudaFree(g->d_mexact_off);
  if (g->d_mexact) cudaFree(g->d_mexact);
  g->d_mexact_off = d_off;
  g->d_mexact = d_b;
  return DMG_OK;
}

int32_t dmg_stacktrace_batch(dmg_graph* g, int32_t k, uint64_t n_traces, const uint32_t* t_off, const uint32_t* cn_off,
                             const uint8_t* cn_bytes, const uint32_t* mn_off, const uint8_t* mn_bytes,
                             uint8_t* out_status, uint32_t* out_vertex, uint32_t* out_method, uint64_t* out_off,
                             uint32_t* out_ids, uint8_t* out_hop, uint64_t cap, uint64_t* needed) {
  if (!g || !needed || !out_off || (n_traces && (!t_off || !cn_off))) {
    set_error("dmg_stacktrace_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (k < 1 || k > DMG_MAX_HOPS) {
    set_error("dmg_stacktrace_batch: k=%d outside 1..%d", k, DMG_MAX_HOPS);
    return DMG_EINVAL;
  }
  if (!g->d_name_off) {
    set_error("dmg_stacktrace_batch: graph was created without names");
    return DMG_EINVAL;
  }
  *needed = 0;
  out_off[0] = 0;
  if (!n_traces) return DMG_OK;
  const uint64_t nf = t_off[n_traces];
  if (nf >= (1ull << 32)) {
    set_error("dmg_stacktrace_batch: too many frames");
    return DMG_EINVAL;
  }
  const uint64_t cb = nf ? cn_off[nf] : 0, mb = (mn_off && nf) ? mn_off[nf] : 0;
  if ((cb && !cn_bytes) || (mb && !mn_bytes) || (nf && (!out_status || !out_vertex))) {
    set_error("dmg_stacktrace_batch: text or per-frame output missing");
    return DMG_EINVAL;
  }
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  const bool with_methods = mn_off && g->d_mexact_off;
  // one staging buffer: cn_off | mn_off | class text | method text ; results: status | vertex | method
  auto pad = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t o_cno = 0, o_mno = pad((nf + 1) * 4), o_cn = o_mno + pad((nf + 1) * 4), o_mn = o_cn + pad(cb + 16),
               in_bytes = o_mn + pad(mb + 16);
  // per-frame results (vertex | method | status) follow the inputs in the same pooled buffer
  const size_t o_vtx = in_bytes, o_meth = o_vtx + pad(nf * 4), o_st = o_meth + pad(nf * 4);
  DMG_TRY(c->q_bytes.reserve(o_st + pad(nf + 16)));
  DMG_TRY(c->q_off.reserve((n_traces + 1) * 4));
  DMG_TRY(c->q_src.reserve(std::max<uint64_t>(nf, 1) * 4));
  uint8_t* in = c->q_bytes.as<uint8_t>();
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, t_off, (n_traces + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (nf) {
    DMG_CUDA(cudaMemcpyAsync(in + o_cno, cn_off, (nf + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    if (cb) DMG_CUDA(cudaMemcpyAsync(in + o_cn, cn_bytes, cb, cudaMemcpyHostToDevice, c->stream));
    if (with_methods) {
      DMG_CUDA(cudaMemcpyAsync(in + o_mno, mn_off, (nf + 1) * 4, cudaMemcpyHostToDevice, c->stream));
      if (mb) DMG_CUDA(cudaMemcpyAsync(in + o_mn, mn_bytes, mb, cudaMemcpyHostToDevice, c->stream));
    }
  }
  c->stats.h2d_bytes += (n_traces + 1) * 4 + 2 * (nf + 1) * 4 + cb + mb;
  StArgs a;
  a.name_off = reinterpret_cast<const unsigned long long*>(g->d_name_off);
  a.names = g->d_names;
  a.full_tbl = g->d_full_tbl;
  a.mask = g->tbl_mask;
  a.V = g->info.n_vertices;
  a.n_parts = g->info.n_parts;
  a.part_off = g->d_part_off;
  a.method_off = g->d_method_off;
  a.mexact_off = with_methods ? g->d_mexact_off : nullptr;
  a.mexact = g->d_mexact;
  a.n_frames = nf;
  a.n_traces = n_traces;
  a.t_off = c->q_off.as<uint32_t>();
  a.cn_off = reinterpret_cast<const uint32_t*>(in + o_cno);
  a.cn = in + o_cn;
  a.mn_off = with_methods ? reinterpret_cast<const uint32_t*>(in + o_mno) : nullptr;
  a.mn = in + o_mn;
  a.vertex = reinterpret_cast<uint32_t*>(in + o_vtx);
  a.method = reinterpret_cast<uint32_t*>(in + o_meth);
  a.status = in + o_st;
  a.q_src = c->q_src.as<uint32_t>();
  if (nf) {
    LaunchScope ls(c, K_SEARCH_FIRST, true, 2);
    st_frames_kernel<<<(unsigned)((nf + 255) / 256), 256, 0, c->stream>>>(a);
    st_sources_kernel<<<(unsigned)((n_traces + 127) / 128), 128, 0, c->stream>>>(a);
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(out_status, a.status, nf, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaMemcpyAsync(out_vertex, a.vertex, nf * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_method) DMG_CUDA(cudaMemcpyAsync(out_method, a.method, nf * 4, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += nf * (out_method ? 9 : 5);
  }
  // neighbour rows: resolveNeighbors (CCS:1558-1578) = mode BOTH, sources excluded, iterated to k
  const uint64_t row_max = g->info.max_part_vertices;
  uint64_t dev_cap = (!out_ids || !out_hop) ? 0 : cap;
  if (row_max && n_traces <= UINT64_MAX / row_max) dev_cap = std::min<uint64_t>(dev_cap, n_traces * row_max);
  DMG_TRY(c->out_off.reserve((n_traces + 1) * 8));
  DMG_TRY(c->out_ids.reserve(dev_cap * 4));
  DMG_TRY(c->out_hop.reserve(dev_cap));
  DMG_TRY(run_khop(c, g, DMG_BOTH, k, 1, n_traces, c->q_off.as<uint32_t>(), c->q_src.as<uint32_t>(),
                   c->out_off.as<uint64_t>(), c->out_ids.as<uint32_t>(), c->out_hop.as<uint8_t>(), dev_cap));
  DMG_CUDA(cudaMemcpyAsync(out_off, c->out_off.p, (n_traces + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  DMG_CUDA(cudaStreamSynchronize(c->stream));
  const uint64_t total = c->h_counters->total;
  c->stats.d2h_bytes += (n_traces + 1) * 8 + sizeof(DevCounters);
  *needed = total;
  int32_t rc = DMG_OK;
  if (total > cap || (total && (!out_ids || !out_hop))) {
    set_error("dmg_stacktrace_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    rc = DMG_ECAP;
  } else if (total) {
    DMG_CUDA(cudaMemcpyAsync(out_ids, c->out_ids.p, total * 4, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaMemcpyAsync(out_hop, c->out_hop.p, total, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += total * 5;
  }
  c->stats.edges_traversed = c->h_counters->edges;
  c->stats.edges_read = c->h_counters->aux[0];
  c->stats.rows_expanded = c->h_counters->rows;
  c->stats.results = total;
  DMG_TRY(c->finish());
  return rc;
}

}  // extern "C"
