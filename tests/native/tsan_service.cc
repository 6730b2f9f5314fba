// ThreadSanitizer driver for the host mirror (csrc/host_graph.cc): GraphService's put / reload / get
// paths (GS:55-136) hammered from several threads.  The device side is replaced by reference-counted
// stand-ins so the run needs no GPU: what is under test is the locking of dmg_service and the
// retain-before-return rule that lets a reader keep a graph a concurrent put() has swapped out.
// Built and run by tests/test_tsan_host_cpu.py with g++ -fsanitize=thread.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dmgraph_host.h"

struct dmg_graph {
  std::atomic<int> refs{1};
  uint32_t n_vertices = 0;
};
struct dmg_engine {
  int unused = 0;
};
static std::atomic<long> g_live{0}, g_created{0};

namespace dmg {
void set_error(const char*, ...) {}
}  // namespace dmg

extern "C" {
int32_t dmg_graph_create(dmg_engine*, const dmg_graph_desc* d, dmg_graph** out) {
  dmg_graph* g = new dmg_graph();
  g->n_vertices = d->n_vertices;
  ++g_live;
  ++g_created;
  *out = g;
  return DMG_OK;
}
void dmg_graph_retain(dmg_graph* g) { g->refs.fetch_add(1, std::memory_order_relaxed); }
void dmg_graph_release(dmg_graph* g) {
  if (g && g->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) {
    --g_live;
    delete g;
  }
}
}

int main() {
  dmg_engine eng;
  dmg_service* svc = nullptr;
  if (dmg_service_new(&eng, &svc) != DMG_OK) return 2;
  // the reload path parses graph_data JSON: take it from the mirror's own writer
  char* json = nullptr;
  {
    dmg_pgraph* pg = dmg_pg_new();
    dmg_pg_add_node(pg, "a.B", "B.java");
    dmg_pg_add_node(pg, "a.C", "C.java");
    dmg_pg_add_dependency(pg, "a.B", "a.C");
    dmg_pg_mark_entry_point(pg, "a.B");
    if (dmg_pg_to_json(pg, &json) != DMG_OK) return 3;
    dmg_pg_free(pg);
  }
  const int kProjects = 4, kWriters = 3, kReaders = 5, kRounds = 300;
  std::atomic<bool> stop{false};
  std::atomic<long> reads{0}, bad{0};
  std::vector<std::thread> th;
  for (int w = 0; w < kWriters; ++w)
    th.emplace_back([&, w] {
      for (int r = 0; r < kRounds; ++r) {
        const int p = (w + r) % kProjects;
        const std::string id = "proj-" + std::to_string(p), name = "name-" + std::to_string(p);
        if (r % 2 == 0) {
          dmg_pgraph* pg = dmg_pg_new();
          const int n = 3 + (r % 5);
          for (int i = 0; i < n; ++i) dmg_pg_add_node(pg, ("co.x.C" + std::to_string(i)).c_str(), "C.java");
          for (int i = 1; i < n; ++i)
            dmg_pg_add_dependency(pg, ("co.x.C" + std::to_string(i - 1)).c_str(), ("co.x.C" + std::to_string(i)).c_str());
          dmg_pg_mark_entry_point(pg, "co.x.C0");
          if (dmg_service_put(svc, id.c_str(), name.c_str(), pg) != DMG_OK) ++bad;
          dmg_pg_free(pg);
        } else {
          if (dmg_service_load_json(svc, id.c_str(), name.c_str(), json) != DMG_OK) ++bad;
        }
      }
    });
  for (int r = 0; r < kReaders; ++r)
    th.emplace_back([&, r] {
      while (!stop.load(std::memory_order_acquire)) {
        const int p = (int)(reads.load(std::memory_order_relaxed) + r) % kProjects;
        dmg_graph* g = (r & 1) ? dmg_service_get(svc, ("proj-" + std::to_string(p)).c_str())
                               : dmg_service_get_by_name(svc, ("name-" + std::to_string(p)).c_str());
        if (g) {
          if (g->n_vertices < 2 || g->n_vertices > 8) ++bad;   // a freed or torn handle would show here (and to TSAN)
          dmg_graph_release(g);
        }
        (void)dmg_service_size(svc);
        ++reads;
      }
    });
  for (int w = 0; w < kWriters; ++w) th[w].join();
  stop.store(true, std::memory_order_release);
  for (size_t i = kWriters; i < th.size(); ++i) th[i].join();
  const uint32_t size = dmg_service_size(svc);
  dmg_service_free(svc);
  dmg_free_string(json);
  std::printf("created=%ld live_after_free=%ld reads=%ld size=%u bad=%ld\n", g_created.load(), g_live.load(), reads.load(),
              size, bad.load());
  return (g_live.load() == 0 && bad.load() == 0 && size == (uint32_t)kProjects) ? 0 : 1;
}
