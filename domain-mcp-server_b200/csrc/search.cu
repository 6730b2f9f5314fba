// S1 / S2 of SURVEY.md 8a: name scans over the packed, pre-lowered identifier arena.
//   S1  search_project        (CCS:1492-1530): every identifier containing the query, nodes order
//   S2  resolveClassName scans (GQS:434-453): first simple-name equality / first containment
//
// v1 kernel: one CTA per query (tickets in query order).  The CTA walks the names of the scan
// range in order, THREADS names at a time; a thread tests its name against the query held in
// shared memory.  Hits of one chunk are compacted in id order (ballot/popc + block scan), so a
// row comes out ascending without a sort.  The row's place in the packed output comes from
// the look-back allocator: the CTA first counts the row, publishes the count, then walks the
// names a second time to write it.
#include <algorithm>

#include "common.cuh"

namespace dmg {

constexpr int SE_THREADS = 256;
constexpr int SE_QMAX = 1024;   // longest query kept in shared memory; longer ones read global

__device__ __forceinline__ bool name_contains(const uint8_t* __restrict__ nm, uint32_t nn,
                                              const uint8_t* __restrict__ pat, uint32_t pn) {
  if (pn == 0) return true;   // "".contains -> every identifier
  if (pn > nn) return false;
  const uint8_t p0 = pat[0];
  for (uint32_t s = 0; s + pn <= nn; ++s) {
    if (nm[s] != p0) continue;
    uint32_t j = 1;
    while (j < pn && nm[s + j] == pat[j]) ++j;
    if (j == pn) return true;
  }
  return false;
}

__device__ __forceinline__ bool simple_name_equals(const uint8_t* __restrict__ nm, uint32_t nn,
                                                   const uint8_t* __restrict__ pat, uint32_t pn) {
  // extractSimpleName: text after the last '.', whole id if there is none (GQS:458-461)
  uint32_t st = 0;
  for (uint32_t i = nn; i > 0; --i)
    if (nm[i - 1] == '.') {
      st = i;
      break;
    }
  if (nn - st != pn) return false;
  for (uint32_t j = 0; j < pn; ++j)
    if (nm[st + j] != pat[j]) return false;
  return true;
}

__global__ void __launch_bounds__(SE_THREADS) search_kernel(const unsigned long long* __restrict__ name_off,
                                                            const uint8_t* __restrict__ lower, uint32_t v_lo,
                                                            uint32_t v_hi, int kind, unsigned long long nq,
                                                            const uint32_t* __restrict__ q_off,
                                                            const uint8_t* __restrict__ q_bytes,
                                                            unsigned long long* out_off, uint32_t* out_ids,
                                                            unsigned long long cap, uint64_t* status,
                                                            DevCounters* ctr) {
  constexpr int THREADS = SE_THREADS, NW = THREADS / 32;
  __shared__ uint8_t s_pat[SE_QMAX];
  __shared__ unsigned long long s_q, s_base;
  __shared__ uint32_t s_w[NW + 1];
  __shared__ uint32_t s_first;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  while (true) {
    if (tid == 0) s_q = atomicAdd(&ctr->ticket, 1ull);
    __syncthreads();
    const unsigned long long q = s_q;
    if (q >= nq) break;
    const uint32_t pb = q_off[q], pn = q_off[q + 1] - pb;
    const uint8_t* pat = q_bytes + pb;
    if (pn <= SE_QMAX) {
      for (uint32_t i = tid; i < pn; i += THREADS) s_pat[i] = pat[i];
      pat = s_pat;
    }
    if (tid == 0) s_first = 0xFFFFFFFFu;
    __syncthreads();
    unsigned long long row_total = 0;
    // pass 0 counts (or finds the first hit), pass 1 writes
    for (int pass = 0; pass < 2; ++pass) {
      unsigned long long run = 0;
      for (uint32_t c0 = v_lo; c0 < v_hi; c0 += THREADS) {
        const uint32_t v = c0 + tid;
        bool hit = false;
        if (v < v_hi) {
          const unsigned long long b = name_off[v];
          const uint32_t nn = (uint32_t)(name_off[v + 1] - b);
          hit = (kind == DMG_SEARCH_SIMPLE_EQ_FIRST) ? simple_name_equals(lower + b, nn, pat, pn)
                                                     : name_contains(lower + b, nn, pat, pn);
        }
        const unsigned m = __ballot_sync(FULL, hit);
        if (kind != DMG_SEARCH_CONTAINS_ALL) {
          if (m && lane == 0) atomicMin(&s_first, c0 + warp * 32 + (__ffs(m) - 1));
          __syncthreads();
          const uint32_t f = s_first;   // read between two barriers: uniform for the whole CTA
          __syncthreads();
          if (f != 0xFFFFFFFFu) break;
          continue;
        }
        if (lane == 0) s_w[warp] = __popc(m);
        __syncthreads();
        if (warp == 0) {
          const uint32_t w = lane < NW ? s_w[lane] : 0;
          const uint32_t iw = warp_incl_scan_u32(w);
          if (lane < NW) s_w[lane] = iw - w;
          if (lane == 31) s_w[NW] = iw;
        }
        __syncthreads();
        if (pass == 1 && hit) {
          const unsigned long long pos = s_base + run + s_w[warp] + __popc(m & lanemask_lt());
          if (pos < cap) out_ids[pos] = v;
        }
        run += s_w[NW];
        __syncthreads();
      }
      if (pass == 0) {
        if (kind != DMG_SEARCH_CONTAINS_ALL) {
          __syncthreads();
          run = (s_first != 0xFFFFFFFFu) ? 1 : 0;
        }
        row_total = run;
        if (warp == 0) {
          const unsigned long long excl = lookback_exclusive(status, q, row_total);
          if (lane == 0) {
            s_base = excl;
            out_off[q] = excl;
            if (q == nq - 1) {
              out_off[nq] = excl + row_total;
              ctr->total = excl + row_total;
            }
          }
        }
        __syncthreads();
        if (kind != DMG_SEARCH_CONTAINS_ALL) {
          if (tid == 0 && row_total && s_base < cap) out_ids[s_base] = s_first;
          break;
        }
        if (row_total == 0) break;
      }
    }
    __syncthreads();
  }
}

}  // namespace dmg

namespace dmg {
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
  DMG_TRY(c->q_off.reserve((nq + 1) * 4));
  DMG_TRY(c->q_bytes.reserve(qbytes + 16));
  DMG_TRY(c->out_off.reserve((nq + 1) * 8));
  DMG_TRY(c->status.reserve(nq * 8));
  // sizing pass first when the worst-case buffer would be unreasonable
  const bool one_pass = dev_cap * 4 <= std::max<uint64_t>(512ull << 20, c->out_ids.cap);
  DMG_CUDA(cudaMemcpyAsync(c->q_off.p, q_off, (nq + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (qbytes) DMG_CUDA(cudaMemcpyAsync(c->q_bytes.p, q_bytes, qbytes, cudaMemcpyHostToDevice, c->stream));
  c->stats.h2d_bytes += (nq + 1) * 4 + qbytes;
  const unsigned grid = (unsigned)std::min<uint64_t>(nq, (uint64_t)g->eng->sm_count * 8);
  for (int pass = one_pass ? 1 : 0; pass < 2; ++pass) {
    const uint64_t use_cap = pass == 0 ? 0 : (one_pass ? dev_cap : std::min<uint64_t>(dev_cap, total));
    if (use_cap) DMG_TRY(c->out_ids.reserve(use_cap * 4));
    DMG_CUDA(cudaMemsetAsync(c->status.p, 0, nq * 8, c->stream));
    DMG_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(DevCounters), c->stream));
    {
      LaunchScope ls(c, K_SEARCH_SCAN);
      search_kernel<<<grid, SE_THREADS, 0, c->stream>>>(
          reinterpret_cast<const unsigned long long*>(g->d_name_off), g->d_lower, v_lo, v_hi, kind, nq,
          c->q_off.as<uint32_t>(), c->q_bytes.as<uint8_t>(), c->out_off.as<unsigned long long>(),
          c->out_ids.as<uint32_t>(), use_cap, c->status.as<uint64_t>(), c->d_counters);
    }
    DMG_CUDA(cudaGetLastError());
    DMG_CUDA(cudaMemcpyAsync(c->h_counters, c->d_counters, sizeof(DevCounters), cudaMemcpyDeviceToHost, c->stream));
    DMG_CUDA(cudaStreamSynchronize(c->stream));
    total = c->h_counters->total;
    if (pass == 0 && (total > (out_ids ? cap : 0) || total == 0)) break;
  }
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
