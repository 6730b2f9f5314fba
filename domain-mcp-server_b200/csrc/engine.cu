// Engine, call contexts, graph upload (GraphService lifecycle, GS:34-122) and small utilities.
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <new>
#include <thread>

#include "common.cuh"

// DMG_UPLOAD_TIMING=1: phase times of an upload on stderr (diagnostics only; nothing is printed otherwise)
namespace {
struct UploadClock {
  bool on;
  std::chrono::steady_clock::time_point t0;
  UploadClock() : on(getenv("DMG_UPLOAD_TIMING") != nullptr), t0(std::chrono::steady_clock::now()) {}
  void tick(const char* what) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[dmg upload] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};
}  // namespace


namespace dmg {

static thread_local char g_err[512] = "";
static thread_local dmg_call_stats g_last_stats{};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int32_t cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  const char* base = strrchr(file, '/');
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), base ? base + 1 : file, line, what);
  if (e == cudaErrorMemoryAllocation) {
    cudaGetLastError();
    return DMG_ENOMEM;
  }
  return (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? DMG_ENODEV : DMG_ECUDA;
}

void publish_stats(const dmg_call_stats& s) { g_last_stats = s; }

const char* const kKernelNames[K_COUNT] = {
    "khop_small", "khop_global", "khop_emit", "scan_counts", "adjacency_gather", "entry_flags", "closure_init", "closure_topdown",
    "closure_bottomup", "order_degrees", "order_minslot", "order_claim", "order_tail", "search_scan",
    "search_first", "search_emit", "khop_windows", "query_rows", "misc", "khop1_rows", "khop_large", "search_replicate"};

int32_t DevBuf::reserve(size_t bytes) {
  if (bytes <= cap && p) return DMG_OK;
  if (p) {
    cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  size_t want = std::max<size_t>(bytes, 256);
  want = (want + 255) & ~size_t(255);
  DMG_CUDA(cudaMalloc(&p, want));
  cap = want;
  return DMG_OK;
}
void DevBuf::release() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}

int32_t CallCtx::init(Engine* e) {
  eng = e;
  DMG_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  DMG_CUDA(cudaEventCreate(&ev_begin));
  DMG_CUDA(cudaEventCreate(&ev_end));
  DMG_CUDA(cudaMalloc((void**)&d_counters, sizeof(DevCounters)));
  DMG_CUDA(cudaMallocHost((void**)&h_counters, sizeof(DevCounters)));
  return DMG_OK;
}
void CallCtx::destroy() {
  for (DevBuf* b : {&status, &scratch, &q_off, &q_src, &q_bytes, &out_off, &out_ids, &out_hop, &tmp0, &tmp1, &tmp2, &tmp3})
    b->release();
  for (auto ev : ev_pool) cudaEventDestroy(ev);
  for (auto& p : prof) {
    cudaEventDestroy(p.a);
    cudaEventDestroy(p.b);
  }
  if (d_counters) cudaFree(d_counters);
  if (h_counters) cudaFreeHost(h_counters);
  if (h_pin) cudaFreeHost(h_pin);
  if (ev_begin) cudaEventDestroy(ev_begin);
  if (ev_end) cudaEventDestroy(ev_end);
  if (stream) cudaStreamDestroy(stream);
}
int32_t CallCtx::reserve_pinned(size_t bytes) {
  if (bytes <= h_pin_cap && h_pin) return DMG_OK;
  if (h_pin) cudaFreeHost(h_pin);
  h_pin = nullptr;
  h_pin_cap = 0;
  DMG_CUDA(cudaMallocHost(&h_pin, std::max<size_t>(bytes, 4096)));
  h_pin_cap = std::max<size_t>(bytes, 4096);
  return DMG_OK;
}
cudaEvent_t CallCtx::get_event() {
  if (!ev_pool.empty()) {
    cudaEvent_t e = ev_pool.back();
    ev_pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
bool nvtx_enabled() {
  static const bool on = [] {
    const char* s = getenv("DMG_NVTX");
    return s && s[0] == '1';
  }();
  return on;
}
void nvtx_push(const char* name) { nvtxRangePushA(name); }
void nvtx_pop() { nvtxRangePop(); }

int32_t CallCtx::begin() {
  if (nvtx_enabled()) nvtx_push("dmg_call");
  memset(&stats, 0, sizeof(stats));
  DMG_CUDA(cudaSetDevice(eng->device));
  DMG_CUDA(cudaEventRecord(ev_begin, stream));
  return DMG_OK;
}
int32_t CallCtx::finish() {
  DMG_CUDA(cudaEventRecord(ev_end, stream));
  DMG_CUDA(cudaStreamSynchronize(stream));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ev_begin, ev_end);
  stats.device_ms = ms;
  if (!prof.empty()) {
    std::lock_guard<std::mutex> lk(eng->prof_mu);
    for (auto& p : prof) {
      float t = 0.f;
      if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) {
        eng->prof_launches[p.kid]++;
        eng->prof_ms[p.kid] += t;
      }
      ev_pool.push_back(p.a);
      ev_pool.push_back(p.b);
    }
    prof.clear();
  }
  publish_stats(stats);
  if (nvtx_enabled()) nvtx_pop();
  return DMG_OK;
}

CallCtx* Engine::acquire() {
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!free_ctx.empty()) {
      CallCtx* c = free_ctx.back();
      free_ctx.pop_back();
      return c;
    }
  }
  cudaSetDevice(device);
  CallCtx* c = new (std::nothrow) CallCtx();
  if (!c) return nullptr;
  if (c->init(this) != DMG_OK) {
    c->destroy();
    delete c;
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(mu);
  all_ctx.push_back(c);
  return c;
}
void Engine::release(CallCtx* c) {
  std::lock_guard<std::mutex> lk(mu);
  free_ctx.push_back(c);
}

void* Engine::pool_take(size_t bytes, size_t* got) {
  bytes = std::max<size_t>((bytes + 255) & ~size_t(255), 256);
  {
    std::lock_guard<std::mutex> lk(pool_mu);
    size_t best = pool.size();
    for (size_t i = 0; i < pool.size(); ++i)   // smallest block that is large enough and not wastefully so
      if (pool[i].second >= bytes && pool[i].second <= 4 * bytes + (1u << 20) &&
          (best == pool.size() || pool[i].second < pool[best].second))
        best = i;
    if (best != pool.size()) {
      void* p = pool[best].first;
      *got = pool[best].second;
      pool_bytes -= pool[best].second;
      pool.erase(pool.begin() + (long)best);
      return p;
    }
  }
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) {
    cudaGetLastError();
    pool_clear();   // the kept blocks may be what is in the way
    if (cudaMalloc(&p, bytes) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
  }
  *got = bytes;
  return p;
}
void Engine::pool_give(void* p, size_t bytes) {
  if (!p) return;
  std::vector<void*> drop;
  {
    std::lock_guard<std::mutex> lk(pool_mu);
    if (bytes > pool_limit) {
      drop.push_back(p);
    } else {
      pool.emplace_back(p, bytes);
      pool_bytes += bytes;
      while (pool.size() > 12 || pool_bytes > pool_limit) {   // oldest first
        drop.push_back(pool.front().first);
        pool_bytes -= pool.front().second;
        pool.erase(pool.begin());
      }
    }
  }
  for (void* d : drop) cudaFree(d);
}
void Engine::pool_clear() {
  std::vector<std::pair<void*, size_t>> all;
  {
    std::lock_guard<std::mutex> lk(pool_mu);
    all.swap(pool);
    pool_bytes = 0;
  }
  for (auto& b : all) cudaFree(b.first);
}

__global__ void flush_kernel(uint4* p, size_t n, uint32_t salt) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = make_uint4(salt, (uint32_t)i, salt ^ 0x5bd1e995u, 0u);
}

}  // namespace dmg

using namespace dmg;

// ---- graph upload ----------------------------------------------------------------------------
static void free_graph(dmg_graph* g) {
  cudaSetDevice(g->eng->device);
  for (void* p : {(void*)g->d_out_ptr, (void*)g->d_out_idx, (void*)g->d_in_ptr, (void*)g->d_in_idx,
                  (void*)g->d_und_ptr, (void*)g->d_und_idx, (void*)g->d_part_off, (void*)g->d_entry, (void*)g->d_entry_by_part, (void*)g->d_entry_bits, (void*)g->d_vmeta, (void*)g->d_name_off,
                  (void*)g->d_lower, (void*)g->d_names, (void*)g->d_full_tbl, (void*)g->d_simple_tbl, (void*)g->d_simple_next, (void*)g->d_method_off, (void*)g->d_mname_off, (void*)g->d_mname_lower, (void*)g->d_mexact_off, (void*)g->d_mexact})
    if (p) cudaFree(p);
  for (auto& el : g->ell)
    for (void* p : {(void*)el.d_proj_slice, (void*)el.d_slice_off, (void*)el.d_vrow, (void*)el.d_data})
      if (p) cudaFree(p);
  delete g;
}

template <class T>
static int32_t upload(T** d, const T* h, size_t n, uint64_t* bytes) {
  size_t b = std::max<size_t>(n * sizeof(T), 16);
  DMG_CUDA(cudaMalloc((void**)d, b));
  if (n) DMG_CUDA(cudaMemcpy(*d, h, n * sizeof(T), cudaMemcpyHostToDevice));
  *bytes += b;
  return DMG_OK;
}

// Sliced-ELL form of one row set, per project (khop_bits.cuh reads it): rows longer than ELL_ROW_CAP
// are cut into virtual rows, the virtual rows of a project are sorted by length (longest first) and
// packed 32 to a slice; inside a slice lane l owns virtual row l and step j holds the j-th id of
// every lane, four steps packed into one uint2 per lane -- so a warp reads a slice with coalesced
// 8-byte loads and every lane ORs its own row without any cross-lane step.  Ids are relative to the
// project base (u16: projects here have <= 4096 vertices); short rows are padded with `pad_id`, the
// index of an all-zero word the kernel keeps behind each set.  fwd_ptr gives the FORWARD degree of
// the row's vertex for the TE count (the pull reads reverse rows).
constexpr uint32_t ELL_ROW_CAP = 256;
namespace {
struct EllPart {   // one project's slices
  std::vector<uint32_t> slice_len;   // steps (multiple of 4) of each slice
  std::vector<uint32_t> vrow;
  std::vector<uint16_t> data;
};
struct EllVR { uint32_t len, v, start; };

void build_ell_project(EllPart& out, const std::vector<uint32_t>& ptr, const std::vector<uint32_t>& idx,
                       const std::vector<uint32_t>& fwd_ptr, uint32_t lo, uint32_t hi, uint32_t pad_id) {
  std::vector<EllVR> vr;
  for (uint32_t v = lo; v < hi; ++v)
    for (uint32_t b = ptr[v]; b < ptr[v + 1]; b += ELL_ROW_CAP) vr.push_back({std::min(ELL_ROW_CAP, ptr[v + 1] - b), v, b});
  std::stable_sort(vr.begin(), vr.end(), [](const EllVR& a, const EllVR& b) { return a.len > b.len; });
  while (vr.size() % 32) vr.push_back({0u, 0xFFFFFFFFu, 0u});
  std::vector<uint16_t> bucket[8][8];   // [lane of the quarter][id % 8], reused
  for (size_t s0 = 0; s0 < vr.size(); s0 += 32) {
    const uint32_t L = (vr[s0].len + 3u) & ~3u;   // longest row of the slice, padded to whole uint2 steps
    const size_t base = out.data.size();
    out.data.resize(base + (size_t)L * 32, (uint16_t)pad_id);
    out.slice_len.push_back(L);
    for (uint32_t l = 0; l < 32; ++l) {
      const EllVR& r = vr[s0 + l];
      if (r.v == 0xFFFFFFFFu) {
        out.vrow.push_back(0xFFFFu);
        continue;
      }
      const uint32_t deg = std::min<uint32_t>(fwd_ptr[r.v + 1] - fwd_ptr[r.v], 0xFFFFu);
      out.vrow.push_back((deg << 16) | (r.v - lo));
    }
    // Order inside a row is free (the kernel ORs), so it is chosen against shared-memory bank
    // conflicts: a 16-byte set word of vertex id sits in bank group id % 8, and the eight lanes of a
    // quarter warp are served together -- at every step each quarter gets eight ids of different
    // groups while the rows still offer them.  Padding reads one of the eight zero words behind the
    // set, again of an unused group.  (ncu, config 2: 23.7 M conflict cycles per launch before, 11.5 M after.)
    for (uint32_t q8 = 0; q8 < 32; q8 += 8) {
      for (uint32_t l = 0; l < 8; ++l) {
        for (int b = 0; b < 8; ++b) bucket[l][b].clear();
        const EllVR& r = vr[s0 + q8 + l];
        if (r.v == 0xFFFFFFFFu) continue;
        for (uint32_t j = 0; j < r.len; ++j) {
          const uint16_t id = (uint16_t)(idx[r.start + j] - lo);
          bucket[l][id & 7].push_back(id);
        }
      }
      for (uint32_t j = 0; j < L; ++j) {
        uint32_t used = 0;
        int choice[8], order[8], nopt[8];
        for (int l = 0; l < 8; ++l) {   // lanes with the fewest groups left choose first
          order[l] = l;
          nopt[l] = 0;
          for (int b = 0; b < 8; ++b) nopt[l] += !bucket[l][b].empty();
        }
        std::sort(order, order + 8, [&](int x, int y) { return nopt[x] < nopt[y]; });
        for (int oi = 0; oi < 8; ++oi) {
          const int l = order[oi];
          int best = -1, best_any = -1;
          for (int b = 0; b < 8; ++b) {
            if (bucket[l][b].empty()) continue;
            if (best_any < 0 || bucket[l][b].size() > bucket[l][best_any].size()) best_any = b;
            if (!(used & (1u << b)) && (best < 0 || bucket[l][b].size() > bucket[l][best].size())) best = b;
          }
          choice[l] = best >= 0 ? best : best_any;
          if (choice[l] >= 0) used |= 1u << choice[l];
        }
        for (int l = 0; l < 8; ++l) {
          uint16_t id;
          if (choice[l] >= 0) {
            id = bucket[l][choice[l]].back();
            bucket[l][choice[l]].pop_back();
          } else {
            int b = 0;
            while (b < 7 && (used & (1u << b))) ++b;
            id = (uint16_t)(pad_id + b);   // words pad_id .. pad_id + 7 of every set are zero
          }
          out.data[base + ((size_t)(j >> 2) * 32 + q8 + l) * 4 + (j & 3)] = id;
        }
      }
    }
  }
}
}  // namespace

static int32_t build_ell(dmg_graph::Ell* out, const std::vector<uint32_t>& ptr, const std::vector<uint32_t>& idx,
                         const std::vector<uint32_t>& fwd_ptr, const std::vector<uint32_t>& part, uint32_t pad_id,
                         uint64_t* bytes) {
  const size_t P = part.size() - 1;
  std::vector<EllPart> parts(P);
  {   // projects are independent: built on the host's threads, stitched in project order
    const unsigned T = (unsigned)std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), P));
    std::atomic<size_t> next{0};
    std::atomic<bool> failed{false};
    auto work = [&] {
      try {
        for (size_t p; (p = next.fetch_add(1)) < P;) build_ell_project(parts[p], ptr, idx, fwd_ptr, part[p], part[p + 1], pad_id);
      } catch (const std::exception&) {
        failed = true;
      }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < T; ++t) th.emplace_back(work);
    work();
    for (auto& x : th) x.join();
    if (failed) {
      set_error("graph upload: host out of memory");
      return DMG_ENOMEM;
    }
  }
  std::vector<uint32_t> proj_slice(P + 1, 0), slice_off(1, 0), vrow;
  std::vector<uint16_t> data;
  size_t n_data = 0, n_vrow = 0;
  for (auto& ep : parts) {
    n_data += ep.data.size();
    n_vrow += ep.vrow.size();
  }
  data.reserve(std::max<size_t>(n_data, 4));
  vrow.reserve(n_vrow);
  for (size_t p = 0; p < P; ++p) {
    EllPart& ep = parts[p];
    for (uint32_t L : ep.slice_len) slice_off.push_back(slice_off.back() + L * 32 / 4);   // uint2 units
    vrow.insert(vrow.end(), ep.vrow.begin(), ep.vrow.end());
    data.insert(data.end(), ep.data.begin(), ep.data.end());
    proj_slice[p + 1] = (uint32_t)(slice_off.size() - 1);
    EllPart().data.swap(ep.data);
  }
  if (data.empty()) data.resize(4, (uint16_t)pad_id);
  out->entries = data.size();
  int32_t st;
  if ((st = upload(&out->d_proj_slice, proj_slice.data(), proj_slice.size(), bytes)) != DMG_OK ||
      (st = upload(&out->d_slice_off, slice_off.data(), slice_off.size(), bytes)) != DMG_OK ||
      (st = upload(&out->d_vrow, vrow.data(), vrow.size(), bytes)) != DMG_OK ||
      (st = upload(reinterpret_cast<uint16_t**>(&out->d_data), data.data(), data.size(), bytes)) != DMG_OK)
    return st;
  return DMG_OK;
}

// Second half of every upload path (edge list or snapshot): derives the resolve rows, the entry
// bitmap and the name tables from the CSR and copies everything to the device.  Takes ownership
// of `g` (freed on failure).
static int32_t finish_upload_impl(dmg_engine* e, dmg_graph* g, uint32_t V, const std::vector<uint32_t>& out_ptr,
                                  const std::vector<uint32_t>& out_idx, const std::vector<uint32_t>& in_ptr,
                                  const std::vector<uint32_t>& in_idx, const std::vector<uint32_t>& part,
                                  const uint64_t* name_off, const uint8_t* name_bytes, const uint8_t* lower_bytes,
                                  uint64_t dropped, uint64_t dups, dmg_graph** out, bool* g_freed) {
  struct Desc {
    const uint64_t* name_off;
    const uint8_t* name_bytes;
    const uint8_t* lower_bytes;
  } dd{name_off, name_bytes, lower_bytes};
  const Desc* d = &dd;
  g->h_part_off = part;
  uint32_t max_part = 0;
  for (size_t i = 1; i < part.size(); ++i) max_part = std::max(max_part, part[i] - part[i - 1]);

  // ---- upload ---------------------------------------------------------------------------------
  cudaSetDevice(e->device);
  uint64_t bytes = 0;
  int32_t st = DMG_OK;
  std::vector<uint32_t> ebits((V + 31) / 32 + 1, 0);
  for (uint32_t v : g->h_entry) ebits[v >> 5] |= 1u << (v & 31);
  std::vector<uint8_t> vmeta(std::max<uint32_t>(V, 1), 0);   // class-type codes arrive with dmg_graph_set_class_types
  for (uint32_t v : g->h_entry) vmeta[v] = 0x80u;
  // entry points grouped by project (stable, so each project keeps markAsEntryPoint order, PG:190-197):
  // executeEntrypoints (GQS:152-191) lists ONE project's graph
  const size_t P = part.size() - 1;
  std::vector<uint32_t> entry_by_part(g->h_entry.size());
  g->h_entry_part_off.assign(P + 1, 0);
  {
    auto part_of = [&](uint32_t v) { return (size_t)(std::upper_bound(part.begin(), part.end(), v) - part.begin() - 1); };
    // zero-width ranges share a boundary: upper_bound lands on the last of them, whose range holds v
    for (uint32_t v : g->h_entry) ++g->h_entry_part_off[std::min(part_of(v), P - 1) + 1];
    for (size_t p = 0; p < P; ++p) g->h_entry_part_off[p + 1] += g->h_entry_part_off[p];
    std::vector<uint32_t> cur(g->h_entry_part_off.begin(), g->h_entry_part_off.end() - 1);
    for (uint32_t v : g->h_entry) entry_by_part[cur[std::min(part_of(v), P - 1)]++] = v;
  }
  // resolve() rows for mode BOTH: the out row followed by the in row of the same vertex, one
  // contiguous range per vertex (no de-duplication: a 2-cycle appears twice, the visited
  // set removes it, PG:366-375)
  UploadClock clk;
  std::vector<uint32_t> und_ptr(V + 1, 0), und_idx;
  try {
    und_idx.resize(2 * out_idx.size());
    for (uint32_t v = 0; v <= V; ++v) und_ptr[v] = out_ptr[v] + in_ptr[v];
    const unsigned T = (unsigned)std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), out_idx.size() / 65536 + 1));
    auto rows = [&](unsigned t) {
      for (uint32_t v = (uint32_t)((uint64_t)V * t / T), v1 = (uint32_t)((uint64_t)V * (t + 1) / T); v < v1; ++v) {
        uint32_t w = und_ptr[v];
        for (uint32_t j = out_ptr[v]; j < out_ptr[v + 1]; ++j) und_idx[w++] = out_idx[j];
        for (uint32_t j = in_ptr[v]; j < in_ptr[v + 1]; ++j) und_idx[w++] = in_idx[j];
      }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < T; ++t) th.emplace_back(rows, t);
    rows(0);
    for (auto& x : th) x.join();
  } catch (const std::bad_alloc&) {
    *g_freed = true;
    delete g;
    set_error("dmg_graph_create: host out of memory");
    return DMG_ENOMEM;
  }
  clk.tick("resolve rows");
  if ((st = upload(&g->d_und_ptr, und_ptr.data(), und_ptr.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_und_idx, und_idx.data(), und_idx.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_out_ptr, out_ptr.data(), out_ptr.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_out_idx, out_idx.data(), out_idx.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_in_ptr, in_ptr.data(), in_ptr.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_in_idx, in_idx.data(), in_idx.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_part_off, part.data(), part.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_entry, g->h_entry.data(), g->h_entry.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_entry_by_part, entry_by_part.data(), entry_by_part.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_entry_bits, ebits.data(), ebits.size(), &bytes)) != DMG_OK ||
      (st = upload(&g->d_vmeta, vmeta.data(), vmeta.size(), &bytes)) != DMG_OK) {
    *g_freed = true;
    free_graph(g);
    return st;
  }
  clk.tick("copies to the device");
  if (max_part <= 4096 && V > 0) {   // MB_MAX_WINDOW of khop_bits.cuh
    const uint32_t pad_id = (std::max<uint32_t>(max_part, 1) + 127u) & ~127u;
    if ((st = build_ell(&g->ell[0], und_ptr, und_idx, und_ptr, part, pad_id, &bytes)) != DMG_OK ||
        (st = build_ell(&g->ell[1], in_ptr, in_idx, out_ptr, part, pad_id, &bytes)) != DMG_OK ||
        (st = build_ell(&g->ell[2], out_ptr, out_idx, in_ptr, part, pad_id, &bytes)) != DMG_OK) {
      *g_freed = true;
      free_graph(g);
      return st;
    }
    g->has_ell = true;
  }
  clk.tick("sliced-ELL copies");
  uint64_t nbytes = 0;
  if (d->name_off) {
    nbytes = d->name_off[V];
    std::vector<uint8_t> low(nbytes + 32, 0);   // zero pad so vector loads may overrun
    const uint8_t* srcb = d->lower_bytes ? d->lower_bytes : d->name_bytes;
    if (nbytes) memcpy(low.data(), srcb, nbytes);
    if (!d->lower_bytes)
      for (uint64_t i = 0; i < nbytes; ++i)
        if (low[i] >= 'A' && low[i] <= 'Z') low[i] += 32;   // ASCII fold only
    // name -> id tables: exact identifier (nodes.containsKey, case
    // sensitive) and lower-cased simple name (GQS:434-446); linear probing, load <= 0.5.
    // Simple names repeat (every TS module id ends in its kind, Go ids share "com/..." tails):
    // the table holds only the FIRST vertex of each (project, simple name); later ones hang off
    // it in ascending id order through simple_next, so neither the build nor a lookup walks a
    // probe run of thousands of equal names.
    uint32_t size = 16;
    while (size < 2ull * V + 1) size <<= 1;
    std::vector<uint32_t> full_tbl(size, 0xFFFFFFFFu), simple_tbl(size, 0xFFFFFFFFu);
    std::vector<uint32_t> simple_next(std::max<uint32_t>(V, 1), 0xFFFFFFFFu);
    std::vector<uint32_t> slot_hash(size, 0), tail(V), sb_rel(V);
    std::vector<uint8_t> orig(nbytes + 32, 0);
    if (nbytes) memcpy(orig.data(), d->name_bytes, nbytes);
    uint32_t part = 0;
    for (uint32_t v = 0; v < V; ++v) {
      const uint64_t b = d->name_off[v], e2 = d->name_off[v + 1];
      uint32_t h = name_hash(orig.data() + b, (uint32_t)(e2 - b)) & (size - 1);
      while (full_tbl[h] != 0xFFFFFFFFu) h = (h + 1) & (size - 1);
      full_tbl[h] = v;
      uint64_t sb = b;
      for (uint64_t i = e2; i > b; --i)
        if (low[i - 1] == '.') {   // extractSimpleName: after the last '.' (GQS:458-461)
          sb = i;
          break;
        }
      sb_rel[v] = (uint32_t)(sb - b);
      while (part + 1 < g->h_part_off.size() - 1 && v >= g->h_part_off[part + 1]) ++part;
      const uint32_t part_lo = g->h_part_off[part], len = (uint32_t)(e2 - sb);
      const uint32_t hv = name_hash(low.data() + sb, len);
      bool chained = false;
      for (h = hv & (size - 1); simple_tbl[h] != 0xFFFFFFFFu; h = (h + 1) & (size - 1)) {
        const uint32_t u = simple_tbl[h];
        if (slot_hash[h] != hv || u < part_lo) continue;
        const uint64_t ub = d->name_off[u] + sb_rel[u], ue = d->name_off[u + 1];
        if (ue - ub == len && memcmp(low.data() + ub, low.data() + sb, len) == 0) {
          simple_next[tail[u]] = v;
          tail[u] = v;
          chained = true;
          break;
        }
      }
      if (!chained) {
        simple_tbl[h] = v;
        slot_hash[h] = hv;
        tail[v] = v;
      }
    }
    g->tbl_mask = size - 1;
    if ((st = upload(&g->d_name_off, d->name_off, (size_t)V + 1, &bytes)) != DMG_OK ||
        (st = upload(&g->d_lower, low.data(), low.size(), &bytes)) != DMG_OK ||
        (st = upload(&g->d_names, orig.data(), orig.size(), &bytes)) != DMG_OK ||
        (st = upload(&g->d_full_tbl, full_tbl.data(), full_tbl.size(), &bytes)) != DMG_OK ||
        (st = upload(&g->d_simple_tbl, simple_tbl.data(), simple_tbl.size(), &bytes)) != DMG_OK ||
        (st = upload(&g->d_simple_next, simple_next.data(), simple_next.size(), &bytes)) != DMG_OK) {
      *g_freed = true;
      free_graph(g);
      return st;
    }
  }
  g->info.n_vertices = V;
  g->info.n_entry = (uint32_t)g->h_entry.size();
  g->info.n_parts = (uint32_t)part.size() - 1;
  g->info.max_part_vertices = max_part;
  g->info.n_edges = out_idx.size();
  g->info.dropped_edges = dropped;
  g->info.duplicate_edges = dups;
  g->info.name_bytes = nbytes;
  g->info.device_bytes = bytes;
  *out = g;
  return DMG_OK;
}

// No C++ exception may cross the extern "C" boundary (it would terminate the host JVM): every
// host allocation of the upload (entry bitmap, resolve rows, name tables) is made under this try.
static int32_t finish_upload(dmg_engine* e, dmg_graph* g, uint32_t V, const std::vector<uint32_t>& out_ptr,
                             const std::vector<uint32_t>& out_idx, const std::vector<uint32_t>& in_ptr,
                             const std::vector<uint32_t>& in_idx, const std::vector<uint32_t>& part,
                             const uint64_t* name_off, const uint8_t* name_bytes, const uint8_t* lower_bytes,
                             uint64_t dropped, uint64_t dups, dmg_graph** out) {
  bool g_freed = false;
  try {
    return finish_upload_impl(e, g, V, out_ptr, out_idx, in_ptr, in_idx, part, name_off, name_bytes, lower_bytes,
                              dropped, dups, out, &g_freed);
  } catch (const std::exception&) {   // bad_alloc, length_error
    if (!g_freed) free_graph(g);
    set_error("graph upload: host out of memory");
    return DMG_ENOMEM;
  }
}

// helpers of the binary CSR snapshot
namespace {
struct SnapHeader {
  char magic[8];
  uint32_t V, n_entry, n_parts, has_names;
  uint64_t E, name_bytes, dropped, dups;
};
template <class T>
bool put_dev(FILE* f, const T* d, size_t n) {
  std::vector<T> h(n);
  if (n && cudaMemcpy(h.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost) != cudaSuccess) return false;
  return !n || fwrite(h.data(), sizeof(T), n, f) == n;
}
template <class T>
bool get_vec(FILE* f, std::vector<T>& v, size_t n) {
  v.resize(n);
  return !n || fread(v.data(), sizeof(T), n, f) == n;
}
}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

const char* dmg_last_error(void) { return g_err; }
const char* dmg_version(void) { return "dmgraph 0.1 (sm_100a)"; }

int32_t dmg_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int32_t dmg_engine_create(const dmg_engine_opts* opts, dmg_engine** out) {
  if (!out) {
    set_error("dmg_engine_create: out is NULL");
    return DMG_EINVAL;
  }
  *out = nullptr;
  int n = 0;
  cudaError_t ce = cudaGetDeviceCount(&n);
  if (ce != cudaSuccess || n == 0) {
    cudaGetLastError();
    set_error("no CUDA device: libdmgraph has no CPU fallback (%s)", ce == cudaSuccess ? "0 devices" : cudaGetErrorString(ce));
    return DMG_ENODEV;
  }
  int dev = opts ? opts->device : 0;
  if (dev < 0 || dev >= n) {
    set_error("device ordinal %d out of range (0..%d)", dev, n - 1);
    return DMG_EINVAL;
  }
  DMG_CUDA(cudaSetDevice(dev));
  cudaDeviceProp prop;
  DMG_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major < 10) {
    set_error("device %d is sm_%d%d; libdmgraph is built for sm_100a only", dev, prop.major, prop.minor);
    return DMG_ENODEV;
  }
  dmg_engine* e = new (std::nothrow) dmg_engine();
  if (!e) return DMG_ENOMEM;
  e->device = dev;
  e->sm_count = prop.multiProcessorCount;
  e->smem_optin = prop.sharedMemPerBlockOptin;
  e->l2_bytes = (size_t)prop.l2CacheSize;
  e->pool_limit = (size_t)prop.totalGlobalMem / 4;
  if (opts && opts->scratch_bytes) e->scratch_budget = opts->scratch_bytes;
  *out = e;
  return DMG_OK;
}

void dmg_engine_destroy(dmg_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  for (CallCtx* c : e->all_ctx) {
    c->destroy();
    delete c;
  }
  if (e->flush_buf) cudaFree(e->flush_buf);
  e->pool_clear();
  delete e;
}

// Host threads of one CSR build: all of them when the build runs alone, an equal share when several
// run side by side (dmg_dist_graph_create builds one graph per device concurrently).
namespace {
std::atomic<int> g_builds_running{0};
struct BuildThreads {
  unsigned n;
  BuildThreads() {
    const int others = g_builds_running.fetch_add(1) + 1;
    n = std::max(1u, std::thread::hardware_concurrency() / (unsigned)others);
  }
  ~BuildThreads() { g_builds_running.fetch_sub(1); }
};
}  // namespace

static int32_t graph_create_impl(dmg_engine* e, const dmg_graph_desc* d, dmg_graph** out);
int32_t dmg_graph_create(dmg_engine* e, const dmg_graph_desc* d, dmg_graph** out) {
  try {
    return graph_create_impl(e, d, out);
  } catch (const std::exception&) {   // the vectors sized from V before the build's own try block
    set_error("dmg_graph_create: host out of memory");
    return DMG_ENOMEM;
  }
}
static int32_t graph_create_impl(dmg_engine* e, const dmg_graph_desc* d, dmg_graph** out) {
  if (!e || !d || !out) {
    set_error("dmg_graph_create: NULL argument");
    return DMG_EINVAL;
  }
  *out = nullptr;
  const uint32_t V = d->n_vertices;
  const uint64_t E = d->n_edges;
  if ((E && (!d->edge_src || !d->edge_dst)) || (d->n_entry && !d->entry_ids)) {
    set_error("dmg_graph_create: edge/entry arrays missing");
    return DMG_EINVAL;
  }
  if (E >= 0x7FFFFFF0ull) {
    set_error("dmg_graph_create: %llu edges exceed the 32-bit CSR offset range", (unsigned long long)E);
    return DMG_EINVAL;
  }
  if (d->name_off && !d->name_bytes && d->name_off[V] != 0) {
    set_error("dmg_graph_create: name_off given without name_bytes");
    return DMG_EINVAL;
  }
  // partitions (projects)
  std::vector<uint32_t> part;
  if (d->n_parts && d->part_off) {
    part.assign(d->part_off, d->part_off + d->n_parts + 1);
    bool ok = part.front() == 0 && part.back() == V;
    for (size_t i = 1; ok && i < part.size(); ++i) ok = part[i] >= part[i - 1];
    if (!ok) {
      set_error("dmg_graph_create: part_off must be non-decreasing from 0 to n_vertices");
      return DMG_EINVAL;
    }
  } else {
    part = {0, V};
  }
  std::vector<uint32_t> part_of;   // only needed to validate multi-project uploads
  auto find_part = [&](uint32_t v) {
    return (uint32_t)(std::upper_bound(part.begin(), part.end(), v) - part.begin() - 1);
  };

  // ---- host CSR build, O(V+E) work on all host threads ---------------------------------------------
  // Stable counting sort by source + per-row de-duplication that keeps the FIRST occurrence
  // (LinkedHashSet.add, PG:182).  Threads own contiguous SOURCE ranges of about equal edge counts: a
  // thread streams the whole edge list in order and places the edges of its own rows, so insertion order
  // is kept without per-thread cursors (T sequential reads of the source array cost far less than the
  // cache-missing scatter they feed).  In-rows are built the same way from the kept edges, threads owning
  // TARGET ranges (step 5).
  uint64_t dropped = 0, dups = 0;
  std::vector<uint32_t> out_ptr(V + 1, 0), in_ptr(V + 1, 0), out_idx, in_idx;
  dmg_graph* g = new (std::nothrow) dmg_graph();
  if (!g) return DMG_ENOMEM;
  g->eng = e;
  try {
    UploadClock clk;
    const BuildThreads bt;   // shares the host's threads with builds running beside this one (dmg_dist_graph_create)
    const unsigned T = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(bt.n, E / 65536 + 1));
    std::atomic<bool> failed{false};
    auto run = [&](auto&& body) {   // body(t): thread index
      std::vector<std::thread> th;
      auto safe = [&](unsigned t) {
        try {
          body(t);
        } catch (const std::exception&) {
          failed = true;
        }
      };
      for (unsigned t = 1; t < T; ++t) th.emplace_back(safe, t);
      safe(0);
      for (auto& x : th) x.join();
      if (failed) throw std::bad_alloc();
    };
    const uint32_t* es = d->edge_src;
    const uint32_t* ed = d->edge_dst;
    // 1. valid edges per source (atomic increments: sources are spread, contention is rare)
    std::vector<uint32_t> cnt(V + 1, 0);
    std::vector<uint64_t> t_dropped(T, 0);
    run([&](unsigned t) {
      const uint64_t a = E * t / T, b = E * (t + 1) / T;
      uint64_t dr = 0;
      for (uint64_t i = a; i < b; ++i) {
        if (es[i] >= V || ed[i] >= V) {   // PG:179-181
          ++dr;
          continue;
        }
        __atomic_fetch_add(&cnt[es[i] + 1], 1u, __ATOMIC_RELAXED);
      }
      t_dropped[t] = dr;
    });
    for (uint64_t x : t_dropped) dropped += x;
    clk.tick("count by source");
    for (uint32_t v = 0; v < V; ++v) cnt[v + 1] += cnt[v];
    const uint64_t kept_raw = cnt[V];
    // source ranges of about equal edge counts
    std::vector<uint32_t> lo_src(T + 1, V);
    lo_src[0] = 0;
    for (unsigned t = 1; t < T; ++t)
      lo_src[t] = (uint32_t)(std::lower_bound(cnt.begin(), cnt.end(), (uint32_t)(kept_raw * t / T)) - cnt.begin());
    for (unsigned t = 1; t <= T; ++t) lo_src[t] = std::max(lo_src[t], lo_src[t - 1]);
    lo_src[T] = V;
    // 2. stable placement, 3. de-duplication in place, kept entries per row in `keep`
    std::vector<uint32_t> raw(kept_raw), keep(V + 1, 0);
    std::vector<uint64_t> t_dups(T, 0);
    run([&](unsigned t) {
      const uint32_t s0 = lo_src[t], s1 = lo_src[t + 1];
      if (s0 >= s1) return;
      std::vector<uint32_t> cur(cnt.begin() + s0, cnt.begin() + s1);
      for (uint64_t i = 0; i < E; ++i) {
        const uint32_t s = es[i];
        if (s < s0 || s >= s1 || ed[i] >= V) continue;
        raw[cur[s - s0]++] = ed[i];
      }
      uint64_t du = 0;
      std::vector<uint32_t> table;   // open addressing, for the few long rows
      for (uint32_t s = s0; s < s1; ++s) {
        const uint32_t b = cnt[s], n = cnt[s + 1] - b;
        uint32_t w = 0;
        if (n <= 48) {
          for (uint32_t j = 0; j < n; ++j) {
            const uint32_t x = raw[b + j];
            bool seen = false;
            for (uint32_t q = 0; q < w && !seen; ++q) seen = raw[b + q] == x;
            if (!seen) raw[b + w++] = x;
          }
        } else {
          uint32_t size = 64;
          while (size < 2 * n) size <<= 1;
          table.assign(size, 0xFFFFFFFFu);
          for (uint32_t j = 0; j < n; ++j) {
            const uint32_t x = raw[b + j];
            uint32_t h = (x * 0x9E3779B1u) & (size - 1);
            while (table[h] != 0xFFFFFFFFu && table[h] != x) h = (h + 1) & (size - 1);
            if (table[h] == x) continue;
            table[h] = x;
            raw[b + w++] = x;
          }
        }
        du += n - w;
        keep[s + 1] = w;
      }
      t_dups[t] = du;
    });
    for (uint64_t x : t_dups) dups += x;
    clk.tick("place + de-duplicate");
    for (uint32_t v = 0; v < V; ++v) out_ptr[v + 1] = out_ptr[v] + keep[v + 1];
    out_idx.resize(out_ptr[V]);
    // 4. compact rows, check that no edge leaves its project
    std::atomic<uint64_t> bad_edge{~0ull};
    run([&](unsigned t) {
      const uint32_t s0 = lo_src[t], s1 = lo_src[t + 1];
      for (uint32_t s = s0; s < s1; ++s) {
        const uint32_t n = out_ptr[s + 1] - out_ptr[s];
        if (!n) continue;
        const uint32_t* from = raw.data() + cnt[s];
        uint32_t* to = out_idx.data() + out_ptr[s];
        uint32_t plo = 0, phi = V;
        if (part.size() > 2) {
          const uint32_t p = find_part(s);
          plo = part[p];
          phi = part[p + 1];
        }
        for (uint32_t j = 0; j < n; ++j) {
          const uint32_t x = from[j];
          to[j] = x;
          if (x < plo || x >= phi) {
            uint64_t want = ~0ull;
            bad_edge.compare_exchange_strong(want, ((uint64_t)s << 32) | x);
          }
        }
      }
    });
    if (bad_edge.load() != ~0ull) {
      set_error("dmg_graph_create: edge %u->%u crosses project ranges", (uint32_t)(bad_edge.load() >> 32),
                (uint32_t)bad_edge.load());
      delete g;
      return DMG_EINVAL;
    }
    std::vector<uint32_t>().swap(raw);
    clk.tick("compact");
    // 5. in-rows.  Import graphs have hubs (one class that 800 k others import): shared cursors would put
    // every thread on the same few counters, and a scattered row would need a sort.  Instead the kept edges
    // are bucketed by TARGET range in one pass: thread t (sources lo_src[t] .. lo_src[t+1]) appends its
    // (target, source) pairs to its own segment of each bucket, segments of a bucket lying in thread order --
    // so a bucket lists its edges in ascending source id.  A bucket is then owned by one thread: private
    // degree counters, private cursors, and every row comes out in ascending source id, its canonical order,
    // by construction.  32 bytes of streamed traffic per edge instead of the kept edge list read once per
    // thread and pass.
    const uint64_t Ek = out_idx.size();
    in_idx.resize(Ek);
    {
      const uint32_t NB = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(1024, (uint64_t)V / 4096 + 1));   // buckets
      const uint32_t bw = (V + NB - 1) / NB ? (V + NB - 1) / NB : 1;                                             // targets per bucket
      std::vector<uint64_t> seg((size_t)T * NB + 1, 0);   // [bucket][thread] segment starts after the prefix
      run([&](unsigned t) {
        std::vector<uint64_t> c(NB, 0);
        for (uint64_t j = out_ptr[lo_src[t]], je = out_ptr[lo_src[t + 1]]; j < je; ++j) ++c[out_idx[j] / bw];
        for (uint32_t b = 0; b < NB; ++b) seg[(size_t)b * T + t + 1] = c[b];
      });
      for (size_t i = 1; i < seg.size(); ++i) seg[i] += seg[i - 1];
      std::vector<uint32_t> px(Ek), ps(Ek);
      run([&](unsigned t) {
        std::vector<uint64_t> cur(NB);
        for (uint32_t b = 0; b < NB; ++b) cur[b] = seg[(size_t)b * T + t];
        for (uint32_t sv = lo_src[t], se = lo_src[t + 1]; sv < se; ++sv)
          for (uint32_t j = out_ptr[sv], je = out_ptr[sv + 1]; j < je; ++j) {
            const uint32_t x = out_idx[j];
            const uint64_t at = cur[x / bw]++;
            px[at] = x;
            ps[at] = sv;
          }
      });
      std::atomic<uint32_t> next_bucket{0};
      run([&](unsigned) {   // in-degrees: a bucket's targets belong to whoever takes the bucket
        uint32_t* deg = in_ptr.data() + 1;
        for (uint32_t b; (b = next_bucket.fetch_add(1)) < NB;)
          for (uint64_t i = seg[(size_t)b * T], ie = seg[(size_t)(b + 1) * T]; i < ie; ++i) ++deg[px[i]];
      });
      for (uint32_t v = 0; v < V; ++v) in_ptr[v + 1] += in_ptr[v];
      next_bucket = 0;
      run([&](unsigned) {
        std::vector<uint32_t> cur;
        for (uint32_t b; (b = next_bucket.fetch_add(1)) < NB;) {
          const uint32_t d0 = (uint32_t)std::min<uint64_t>(V, (uint64_t)b * bw), d1 = (uint32_t)std::min<uint64_t>(V, (uint64_t)(b + 1) * bw);
          cur.assign(in_ptr.begin() + d0, in_ptr.begin() + d1);
          for (uint64_t i = seg[(size_t)b * T], ie = seg[(size_t)(b + 1) * T]; i < ie; ++i) in_idx[cur[px[i] - d0]++] = ps[i];
        }
      });
    }
    clk.tick("in-rows (bucket + place)");
    // entry points: insertion order, unknown ids ignored, no duplicates (PG:190-197)
    std::vector<uint8_t> is_e(V, 0);
    for (uint32_t i = 0; i < d->n_entry; ++i) {
      const uint32_t v = d->entry_ids[i];
      if (v < V && !is_e[v]) {
        is_e[v] = 1;
        g->h_entry.push_back(v);
      }
    }
  } catch (const std::bad_alloc&) {
    delete g;
    set_error("dmg_graph_create: host out of memory");
    return DMG_ENOMEM;
  }
  return finish_upload(e, g, V, out_ptr, out_idx, in_ptr, in_idx, part, d->name_off, d->name_bytes, d->lower_bytes,
                       dropped, dups, out);
}

// ---- binary CSR snapshot ------------------------------------------------
int32_t dmg_graph_save(dmg_graph* g, const char* path) {
  if (!g || !path) {
    set_error("dmg_graph_save: NULL argument");
    return DMG_EINVAL;
  }
  DMG_CUDA(cudaSetDevice(g->eng->device));
  FILE* f = fopen(path, "wb");
  if (!f) {
    set_error("dmg_graph_save: cannot open %s", path);
    return DMG_EINVAL;
  }
  SnapHeader h;
  memcpy(h.magic, "DMGCSR1", 8);
  h.V = g->info.n_vertices;
  h.n_entry = g->info.n_entry;
  h.n_parts = g->info.n_parts;
  h.has_names = g->d_name_off != nullptr;
  h.E = g->info.n_edges;
  h.name_bytes = g->info.name_bytes;
  h.dropped = g->info.dropped_edges;
  h.dups = g->info.duplicate_edges;
  bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && put_dev(f, g->d_out_ptr, (size_t)h.V + 1) &&
            put_dev(f, g->d_out_idx, h.E) && put_dev(f, g->d_in_ptr, (size_t)h.V + 1) && put_dev(f, g->d_in_idx, h.E) &&
            fwrite(g->h_part_off.data(), 4, g->h_part_off.size(), f) == g->h_part_off.size() &&
            (g->h_entry.empty() || fwrite(g->h_entry.data(), 4, g->h_entry.size(), f) == g->h_entry.size());
  if (ok && h.has_names)
    ok = put_dev(f, g->d_name_off, (size_t)h.V + 1) && put_dev(f, g->d_names, h.name_bytes) &&
         put_dev(f, g->d_lower, h.name_bytes);
  ok = (fclose(f) == 0) && ok;
  if (!ok) {
    set_error("dmg_graph_save: write to %s failed", path);
    return DMG_EINVAL;
  }
  return DMG_OK;
}

// A snapshot is a file somebody else may have truncated, edited or written for another graph:
// it is held to everything dmg_graph_create guarantees (kernels index shared memory with
// `id - project base` unchecked), and no allocation is sized from a header field before the
// file's length has confirmed it.
static int32_t load_snapshot(dmg_engine* e, const char* path, FILE* f, dmg_graph** out) {
  SnapHeader h;
  std::vector<uint32_t> out_ptr, out_idx, in_ptr, in_idx, part, entry;
  std::vector<uint64_t> name_off;
  std::vector<uint8_t> names, lower;
  bool ok = fread(&h, sizeof(h), 1, f) == 1 && memcmp(h.magic, "DMGCSR1", 8) == 0 && h.E < 0x7FFFFFF0ull &&
            h.n_parts >= 1 && h.has_names <= 1;
  if (ok) {
    unsigned long long want = sizeof(h) + 2ull * ((unsigned long long)h.V + 1) * 4 + 2ull * h.E * 4 +
                              ((unsigned long long)h.n_parts + 1) * 4 + (unsigned long long)h.n_entry * 4;
    if (h.has_names) {
      ok = h.name_bytes < (1ull << 40);
      want += ((unsigned long long)h.V + 1) * 8 + 2ull * h.name_bytes;
    }
    long long size = -1;
    if (fseek(f, 0, SEEK_END) == 0) size = ftell(f);
    ok = ok && size >= 0 && (unsigned long long)size == want && fseek(f, (long)sizeof(h), SEEK_SET) == 0;
  }
  ok = ok && get_vec(f, out_ptr, (size_t)h.V + 1) && get_vec(f, out_idx, h.E) && get_vec(f, in_ptr, (size_t)h.V + 1) &&
       get_vec(f, in_idx, h.E) && get_vec(f, part, (size_t)h.n_parts + 1) && get_vec(f, entry, h.n_entry);
  if (ok && h.has_names)
    ok = get_vec(f, name_off, (size_t)h.V + 1) && get_vec(f, names, h.name_bytes) && get_vec(f, lower, h.name_bytes);
  // shape: offsets monotone and in range, project ranges non-decreasing from 0 to V
  ok = ok && out_ptr[0] == 0 && in_ptr[0] == 0 && out_ptr[h.V] == h.E && in_ptr[h.V] == h.E && part.front() == 0 &&
       part.back() == h.V;
  for (uint32_t v = 0; ok && v < h.V; ++v) ok = out_ptr[v] <= out_ptr[v + 1] && in_ptr[v] <= in_ptr[v + 1];
  for (size_t i = 1; ok && i < part.size(); ++i) ok = part[i] >= part[i - 1];
  for (uint64_t i = 0; ok && i < h.E; ++i) ok = out_idx[i] < h.V;
  // every edge stays inside the project of its source (what dmg_graph_create rejects)
  if (ok) {
    size_t p = 0;
    for (uint32_t s = 0; ok && s < h.V; ++s) {
      while (s >= part[p + 1]) ++p;
      for (uint32_t j = out_ptr[s]; ok && j < out_ptr[s + 1]; ++j) ok = out_idx[j] >= part[p] && out_idx[j] < part[p + 1];
    }
  }
  // the reverse CSR must be the transpose of the forward one, rows in ascending source id
  if (ok) {
    std::vector<uint32_t> cur((size_t)h.V + 1, 0);
    for (uint64_t i = 0; i < h.E; ++i) ++cur[out_idx[i] + 1];
    for (uint32_t v = 0; v < h.V; ++v) cur[v + 1] += cur[v];
    ok = memcmp(cur.data(), in_ptr.data(), ((size_t)h.V + 1) * 4) == 0;
    for (uint32_t s = 0; ok && s < h.V; ++s)
      for (uint32_t j = out_ptr[s]; ok && j < out_ptr[s + 1]; ++j) ok = in_idx[cur[out_idx[j]]++] == s;
  }
  // entry points: known ids, no repeats (PG:190-197)
  if (ok) {
    std::vector<uint8_t> seen(h.V, 0);
    for (uint32_t i = 0; ok && i < h.n_entry; ++i) {
      ok = entry[i] < h.V && !seen[entry[i]];
      if (ok) seen[entry[i]] = 1;
    }
  }
  if (ok && h.has_names) {
    ok = name_off[0] == 0 && name_off[h.V] == h.name_bytes;
    for (uint32_t v = 0; ok && v < h.V; ++v) ok = name_off[v] <= name_off[v + 1];
  }
  if (!ok) {
    set_error("dmg_graph_load: %s is not a valid dmgraph snapshot", path);
    return DMG_EINVAL;
  }
  dmg_graph* g = new (std::nothrow) dmg_graph();
  if (!g) return DMG_ENOMEM;
  g->eng = e;
  g->h_entry.swap(entry);   // no allocation: nothing can throw between `new` and finish_upload
  return finish_upload(e, g, h.V, out_ptr, out_idx, in_ptr, in_idx, part, h.has_names ? name_off.data() : nullptr,
                       names.data(), h.has_names ? lower.data() : nullptr, h.dropped, h.dups, out);
}

int32_t dmg_graph_load(dmg_engine* e, const char* path, dmg_graph** out) {
  if (!e || !path || !out) {
    set_error("dmg_graph_load: NULL argument");
    return DMG_EINVAL;
  }
  *out = nullptr;
  FILE* f = fopen(path, "rb");
  if (!f) {
    set_error("dmg_graph_load: cannot open %s", path);
    return DMG_EINVAL;
  }
  int32_t rc;
  try {
    rc = load_snapshot(e, path, f, out);
  } catch (const std::exception&) {
    set_error("dmg_graph_load: host out of memory reading %s", path);
    rc = DMG_ENOMEM;
  }
  fclose(f);
  return rc;
}

void dmg_graph_retain(dmg_graph* g) {
  if (g) g->refs.fetch_add(1);
}
void dmg_graph_release(dmg_graph* g) {
  if (g && g->refs.fetch_sub(1) == 1) free_graph(g);
}
int32_t dmg_graph_info_get(const dmg_graph* g, dmg_graph_info* out) {
  if (!g || !out) {
    set_error("dmg_graph_info_get: NULL argument");
    return DMG_EINVAL;
  }
  *out = g->info;
  return DMG_OK;
}

// ---- stats / profiling ------------------------------------------------------------------------
int32_t dmg_last_call_stats(dmg_engine* e, dmg_call_stats* out) {
  if (!e || !out) return DMG_EINVAL;
  *out = g_last_stats;
  return DMG_OK;
}
int32_t dmg_profile_enable(dmg_engine* e, int32_t on) {
  if (!e) return DMG_EINVAL;
  e->profile.store(on ? 1 : 0);
  return DMG_OK;
}
int32_t dmg_profile_reset(dmg_engine* e) {
  if (!e) return DMG_EINVAL;
  std::lock_guard<std::mutex> lk(e->prof_mu);
  for (int i = 0; i < K_COUNT; ++i) {
    e->prof_launches[i] = 0;
    e->prof_ms[i] = 0;
  }
  return DMG_OK;
}
int32_t dmg_profile_count(dmg_engine*) { return K_COUNT; }
int32_t dmg_profile_get(dmg_engine* e, int32_t i, const char** name, uint64_t* launches, double* ms) {
  if (!e || i < 0 || i >= K_COUNT) return DMG_EINVAL;
  std::lock_guard<std::mutex> lk(e->prof_mu);
  if (name) *name = kKernelNames[i];
  if (launches) *launches = e->prof_launches[i];
  if (ms) *ms = e->prof_ms[i];
  return DMG_OK;
}

// ---- memory helpers ---------------------------------------------------------------------------
int32_t dmg_device_alloc(dmg_engine* e, uint64_t bytes, void** p) {
  if (!e || !p) return DMG_EINVAL;
  DMG_CUDA(cudaSetDevice(e->device));
  DMG_CUDA(cudaMalloc(p, std::max<uint64_t>(bytes, 16)));
  return DMG_OK;
}
int32_t dmg_device_free(dmg_engine* e, void* p) {
  if (!e) return DMG_EINVAL;
  DMG_CUDA(cudaSetDevice(e->device));
  DMG_CUDA(cudaFree(p));
  return DMG_OK;
}
int32_t dmg_copy_to_device(dmg_engine* e, void* d, const void* s, uint64_t bytes) {
  if (!e) return DMG_EINVAL;
  DMG_CUDA(cudaSetDevice(e->device));
  DMG_CUDA(cudaMemcpy(d, s, bytes, cudaMemcpyHostToDevice));
  return DMG_OK;
}
int32_t dmg_copy_to_host(dmg_engine* e, void* d, const void* s, uint64_t bytes) {
  if (!e) return DMG_EINVAL;
  DMG_CUDA(cudaSetDevice(e->device));
  DMG_CUDA(cudaMemcpy(d, s, bytes, cudaMemcpyDeviceToHost));
  return DMG_OK;
}
int32_t dmg_pinned_alloc(dmg_engine* e, uint64_t bytes, void** p) {
  if (!e || !p) return DMG_EINVAL;
  DMG_CUDA(cudaSetDevice(e->device));
  DMG_CUDA(cudaMallocHost(p, std::max<uint64_t>(bytes, 16)));
  return DMG_OK;
}
int32_t dmg_pinned_free(dmg_engine* e, void* p) {
  if (!e) return DMG_EINVAL;
  DMG_CUDA(cudaFreeHost(p));
  return DMG_OK;
}
int32_t dmg_flush_l2(dmg_engine* e) {
  if (!e) return DMG_EINVAL;
  DMG_CUDA(cudaSetDevice(e->device));
  if (!e->flush_buf) {
    e->flush_bytes = std::max<size_t>(2 * e->l2_bytes, 256u << 20);
    DMG_CUDA(cudaMalloc(&e->flush_buf, e->flush_bytes));
  }
  static std::atomic<uint32_t> salt{1};
  flush_kernel<<<e->sm_count * 8, 256>>>((uint4*)e->flush_buf, e->flush_bytes / 16, salt.fetch_add(1));
  DMG_CUDA(cudaGetLastError());
  DMG_CUDA(cudaDeviceSynchronize());
  return DMG_OK;
}

}  // extern "C"
