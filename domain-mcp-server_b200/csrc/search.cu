// S1 / S2 of SURVEY.md 8a: the entry point of the name scans over the packed, pre-lowered
// identifier arena.
//   S1  search_project (CCS:1492-1530) and the containment step of resolveClassName (GQS:449-453):
//       search_multi.cu, one pass over the arena for the whole batch;
//   S2  the simple-name step of resolveClassName (GQS:434-446): a probe of the hash table built
//       at upload (lookup.cu) instead of a scan.
#include <algorithm>
#include <vector>

#include "common.cuh"

namespace dmg {
int32_t run_lookup(CallCtx* c, dmg_graph* g, int kind, uint32_t v_lo, uint32_t v_hi, uint64_t n,
                   const uint32_t* d_t_off, const uint8_t* d_text, uint32_t* d_out);
int32_t run_search_multi(CallCtx* c, dmg_graph* g, int kind, uint32_t v_lo, uint32_t v_hi, uint64_t nq,
                         const uint32_t* q_off, const uint8_t* q_bytes, uint64_t dev_cap, bool sizes_only);
}

using namespace dmg;

extern "C" {

int32_t dmg_search_batch(dmg_graph* g, int32_t kind, uint32_t v_lo, uint32_t v_hi, uint64_t nq, const uint32_t* q_off,
                         const uint8_t* q_bytes, uint64_t* out_off, uint32_t* out_ids, uint64_t cap,
                         uint64_t* needed) {
  if (!g || !out_off || !needed || (nq && !q_off)) {
    set_error("dmg_search_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (kind < DMG_SEARCH_CONTAINS_ALL || kind > DMG_SEARCH_CONTAINS_FIRST) {
    set_error("dmg_search_batch: unknown kind %d", kind);
    return DMG_EINVAL;
  }
  if (!g->d_name_off) {
    set_error("dmg_search_batch: graph was created without names");
    return DMG_EINVAL;
  }
  if (v_hi > g->info.n_vertices) v_hi = g->info.n_vertices;
  if (v_lo > v_hi) v_lo = v_hi;
  *needed = 0;
  out_off[0] = 0;
  if (nq == 0) return DMG_OK;
  constexpr uint64_t SLICE = 16384;   // distinct patterns one pass over the arena handles
  if (nq > SLICE) {
    // larger batches run slice by slice; rows are appended in query order
    uint64_t total = 0;
    int32_t rc = DMG_OK;
    std::vector<uint64_t> off(SLICE + 1);
    for (uint64_t q0 = 0; q0 < nq; q0 += SLICE) {
      const uint64_t n = std::min<uint64_t>(SLICE, nq - q0);
      uint64_t need = 0;
      const uint64_t room = (rc == DMG_OK && out_ids && cap > total) ? cap - total : 0;
      const int32_t st = dmg_search_batch(g, kind, v_lo, v_hi, n, q_off + q0, q_bytes, off.data(),
                                          room ? out_ids + total : nullptr, room, &need);
      if (st != DMG_OK && st != DMG_ECAP) return st;
      if (st == DMG_ECAP) rc = DMG_ECAP;
      for (uint64_t j = 0; j < n; ++j) out_off[q0 + j] = total + off[j];
      total += need;
    }
    out_off[nq] = total;
    *needed = total;
    if (rc == DMG_ECAP)
      set_error("dmg_search_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    return rc;
  }
  const uint64_t qbytes = q_off[nq];
  if (qbytes && !q_bytes) {
    set_error("dmg_search_batch: q_bytes is NULL (a null query is the host's NullPointerException, CCS:1507)");
    return DMG_EINVAL;
  }
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  const uint64_t range = v_hi - v_lo;
  uint64_t dev_cap = out_ids ? cap : 0;
  if (kind != DMG_SEARCH_CONTAINS_ALL) dev_cap = std::min<uint64_t>(dev_cap, nq);
  else if (range == 0 || nq <= UINT64_MAX / std::max<uint64_t>(range, 1)) dev_cap = std::min<uint64_t>(dev_cap, nq * range);
  uint64_t total = 0;
  if (kind != DMG_SEARCH_SIMPLE_EQ_FIRST) {
    // S1 and the containment scan of S2: one pass over the arena for the whole batch
    const uint64_t cap_eff = out_ids ? cap : 0;
    const bool direct = cap_eff && dev_cap * 4 <= std::max<uint64_t>(1ull << 30, c->out_ids.cap);
    DMG_TRY(run_search_multi(c, g, kind, v_lo, v_hi, nq, q_off, q_bytes, direct ? dev_cap : 0, !direct));
    total = c->h_counters->total;
    if (!direct && cap_eff && total && total <= cap_eff) {
      DMG_TRY(run_search_multi(c, g, kind, v_lo, v_hi, nq, q_off, q_bytes, total, false));
      total = c->h_counters->total;
    }
  } else {
    // first vertex whose lower-cased simple name equals the query: hash probes, no scan
    DMG_TRY(c->q_off.reserve((nq + 1) * 4));
    DMG_TRY(c->q_bytes.reserve(qbytes + 16));
    DMG_TRY(c->out_ids.reserve(nq * 4));
    DMG_CUDA(cudaMemcpyAsync(c->q_off.p, q_off, (nq + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    if (qbytes) DMG_CUDA(cudaMemcpyAsync(c->q_bytes.p, q_bytes, qbytes, cudaMemcpyHostToDevice, c->stream));
    c->stats.h2d_bytes += (nq + 1) * 4 + qbytes;
    DMG_TRY(run_lookup(c, g, DMG_LOOKUP_SIMPLE_FIRST, v_lo, v_hi, nq, c->q_off.as<uint32_t>(),
                       c->q_bytes.as<uint8_t>(), c->out_ids.as<uint32_t>()));
    std::vector<uint32_t> found(nq);
    DMG_CUDA(cudaMemcpyAsync(found.data(), c->out_ids.p, nq * 4, cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    c->stats.d2h_bytes += nq * 4;
    for (uint64_t q = 0; q < nq; ++q) {
      out_off[q] = total;
      if (found[q] != 0xFFFFFFFFu) {
        if (out_ids && total < cap) out_ids[total] = found[q];
        ++total;
      }
    }
    out_off[nq] = total;
    *needed = total;
    c->stats.results = total;
    DMG_TRY(c->finish());
    if (total > cap || (total && !out_ids)) {
      set_error("dmg_search_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
      return DMG_ECAP;
    }
    return DMG_OK;
  }
  DMG_CUDA(cudaMemcpyAsync(out_off, c->out_off.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  c->stats.d2h_bytes += (nq + 1) * 8 + sizeof(DevCounters);
  *needed = total;
  int32_t rc = DMG_OK;
  if (total > cap || (total && !out_ids)) {
    set_error("dmg_search_batch: capacity %llu < needed %llu", (unsigned long long)cap, (unsigned long long)total);
    rc = DMG_ECAP;
  } else if (total) {
    DMG_CUDA(cudaMemcpyAsync(out_ids, c->out_ids.p, total * 4, cudaMemcpyDeviceToHost, c->stream));
    c->stats.d2h_bytes += total * 4;
  }
  c->stats.results = total;
  DMG_TRY(c->finish());
  return rc;
}

}  // extern "C"
