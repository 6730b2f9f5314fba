// Name -> id lookups against the hash tables built at upload.
#include <algorithm>

#include "common.cuh"

namespace dmg {

__global__ void __launch_bounds__(256) lookup_kernel(const unsigned long long* __restrict__ name_off,
                                                     const uint8_t* __restrict__ names,
                                                     const uint8_t* __restrict__ lower,
                                                     const uint32_t* __restrict__ tbl, uint32_t mask,
                                                     const uint32_t* __restrict__ simple_next,
                                                     const uint32_t* __restrict__ part_off, uint32_t n_parts, int kind,
                                                     uint32_t v_lo, uint32_t v_hi, unsigned long long n,
                                                     const uint32_t* __restrict__ t_off,
                                                     const uint8_t* __restrict__ text, uint32_t* out) {
  const unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t* q = text + t_off[i];
  const uint32_t qn = t_off[i + 1] - t_off[i];
  uint32_t best = 0xFFFFFFFFu;
  for (uint32_t h = name_hash(q, qn) & mask;; h = (h + 1) & mask) {
    const uint32_t v = tbl[h];
    if (v == 0xFFFFFFFFu) break;
    // the simple table holds chain heads: everything chained behind v has a larger id
    if (v >= v_hi || v >= best || (kind == DMG_LOOKUP_EXACT && v < v_lo)) continue;
    const unsigned long long b = name_off[v], e = name_off[v + 1];
    unsigned long long s = b;
    const uint8_t* arena = names;
    if (kind == DMG_LOOKUP_SIMPLE_FIRST) {
      arena = lower;
      for (unsigned long long j = e; j > b; --j)
        if (lower[j - 1] == '.') {
          s = j;
          break;
        }
    }
    if (e - s != qn) continue;
    uint32_t j = 0;
    while (j < qn && arena[s + j] == q[j]) ++j;
    if (j != qn) continue;
    if (v >= v_lo) {
      best = v;
      if (kind == DMG_LOOKUP_EXACT) break;   // identifiers are unique inside a project
      continue;
    }
    // SIMPLE_FIRST, head before the range: its project may still reach into the range
    uint32_t l = 0, r = n_parts;
    while (r - l > 1) {
      const uint32_t m = (l + r) >> 1;
      if (part_off[m] <= v) l = m; else r = m;
    }
    if (part_off[l + 1] <= v_lo) continue;
    uint32_t w = simple_next[v];
    while (w != 0xFFFFFFFFu && w < v_lo) w = simple_next[w];
    if (w < v_hi && w < best) best = w;   // 0xFFFFFFFF fails both tests
  }
  out[i] = best;
}

// device-buffer form used by dmg_search_batch(SIMPLE_EQ_FIRST) and dmg_lookup_batch
int32_t run_lookup(CallCtx* c, dmg_graph* g, int kind, uint32_t v_lo, uint32_t v_hi, uint64_t n,
                   const uint32_t* d_t_off, const uint8_t* d_text, uint32_t* d_out) {
  if (!n) return DMG_OK;
  LaunchScope ls(c, K_SEARCH_FIRST);
  lookup_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(
      reinterpret_cast<const unsigned long long*>(g->d_name_off), g->d_names, g->d_lower,
      kind == DMG_LOOKUP_EXACT ? g->d_full_tbl : g->d_simple_tbl, g->tbl_mask, g->d_simple_next, g->d_part_off,
      g->info.n_parts, kind, v_lo, v_hi, n, d_t_off, d_text, d_out);
  DMG_CUDA(cudaGetLastError());
  return DMG_OK;
}

}  // namespace dmg

using namespace dmg;

extern "C" int32_t dmg_lookup_batch(dmg_graph* g, int32_t kind, uint32_t v_lo, uint32_t v_hi, uint64_t n,
                                    const uint32_t* t_off, const uint8_t* t_bytes, uint32_t* out_ids) {
  if (!g || (n && (!t_off || !out_ids))) {
    set_error("dmg_lookup_batch: NULL argument");
    return DMG_EINVAL;
  }
  if (kind != DMG_LOOKUP_EXACT && kind != DMG_LOOKUP_SIMPLE_FIRST) {
    set_error("dmg_lookup_batch: unknown kind %d", kind);
    return DMG_EINVAL;
  }
  if (!g->d_name_off) {
    set_error("dmg_lookup_batch: graph was created without names");
    return DMG_EINVAL;
  }
  if (!n) return DMG_OK;
  const uint64_t tb = t_off[n];
  if (tb && !t_bytes) {
    set_error("dmg_lookup_batch: t_bytes is NULL");
    return DMG_EINVAL;
  }
  v_hi = std::min(v_hi, g->info.n_vertices);
  CtxGuard cg(g->eng);
  CallCtx* c = cg.c;
  if (!c) return DMG_ECUDA;
  DMG_TRY(c->begin());
  DMG_TRY(c->q_off.reserve((n + 1) * 4));
  DMG_TRY(c->q_bytes.reserve(tb + 16));
  DMG_TRY(c->out_ids.reserve(n * 4));
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, t_off, (n + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (tb) DMG_CUDA(cudaMemcpyAsync(c->q_bytes.p, t_bytes, tb, cudaMemcpyHostToDevice, c->stream));
  DMG_TRY(run_lookup(c, g, kind, v_lo, v_hi, n, c->q_off.as<uint32_t>(), c->q_bytes.as<uint8_t>(),
                     c->out_ids.as<uint32_t>()));
  DMG_CUDA(cudaMemcpyAsync(out_ids, c->out_ids.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  c->stats.h2d_bytes += (n + 1) * 4 + tb;
  c->stats.d2h_bytes += n * 4;
  c->stats.results = n;
  return c->finish();
}
