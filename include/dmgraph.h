/*
 * dmgraph.h -- C ABI of libdmgraph.so, the B200-resident graph + name-search engine that
 * stands in for the hot path of waabox/domain-mcp-server.
 *
 * The reference has no native seam today: its de-facto interface is the read
 * API of ProjectGraph as consumed by CodeContextService and GraphQueryService, plus the
 * GraphService lifecycle.  Each entry point below names the reference code it replaces
 * (paths relative to src/main/java/co/fanki/domainmcp/ in the reference):
 *
 *   PG  = analysis/domain/ProjectGraph.java
 *   GS  = analysis/domain/GraphService.java
 *   CCS = analysis/application/CodeContextService.java
 *   GQS = analysis/application/GraphQueryService.java
 *
 * Conventions
 *   - Vertex id (uint32) = position of the identifier in the `nodes` sequence the host
 *     supplies (PG:118 LinkedHashMap order).  The host keeps String<->id.
 *   - Every buffer is caller-owned.  Inputs are read-only during the call.  Outputs are
 *     caller-allocated with a capacity; if the result does not fit the call returns
 *     DMG_ECAP, fills out_off[] and *needed, and writes no row data.
 *   - An unknown / out-of-range vertex id in a query yields an empty row, never an error
 *     (PG:300-302, 317-319, 360-362).
 *   - All query entry points may be called concurrently from any number of host threads on
 *     one graph handle (the reference's readers are lock-free, GS:34-38).
 *   - The library never writes to stdout or stderr (MCP stdio transport) and never aborts.
 *   - There is no CPU fallback: without a CUDA device every entry point that needs one
 *     fails with DMG_ENODEV / DMG_ECUDA.
 *   - Pointers named d_* are DEVICE pointers (from dmg_device_alloc); all others are host.
 */
#ifndef DMGRAPH_H
#define DMGRAPH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DMG_OK 0
#define DMG_EINVAL 1 /* IllegalArgumentException analogue (shared/Preconditions.java:41-47) */
#define DMG_ENOMEM 2
#define DMG_ECAP 3   /* output capacity too small; *needed says how much */
#define DMG_ECUDA 4
#define DMG_ENCCL 5
#define DMG_ENODEV 6

/* traversal direction / mode */
#define DMG_OUT 0  /* dependencies  (PG:299-305) */
#define DMG_IN 1   /* dependents    (PG:316-328) */
#define DMG_BOTH 2 /* resolve       (PG:359-378) */

/* search kinds */
#define DMG_SEARCH_CONTAINS_ALL 0   /* CCS:1511-1526 search_project: every hit, nodes order */
#define DMG_SEARCH_SIMPLE_EQ_FIRST 1 /* GQS:434-446 first id whose simple name equals query */
#define DMG_SEARCH_CONTAINS_FIRST 2 /* GQS:449-453 first id containing the query */

#define DMG_MAX_HOPS 254
#define DMG_LEVEL_UNREACHED 0xFFFFFFFFu

typedef struct dmg_engine dmg_engine;
typedef struct dmg_graph dmg_graph;

typedef struct dmg_engine_opts {
  int32_t device;         /* CUDA ordinal this engine runs on */
  int32_t flags;          /* reserved, 0 */
  uint64_t scratch_bytes; /* per-call traversal scratch budget; 0 = default (4 GiB) */
} dmg_engine_opts;

/* Host-side description of one immutable graph (all projects of a GraphService, or one).
 * Replaces: new ProjectGraph() + addNode/addDependency/markAsEntryPoint (PG:160-197) followed
 * by GraphService.put (GS:118-122).  Build rules reproduced at upload:
 *   - an edge is kept only if both endpoints are < n_vertices (PG:179-181), else dropped
 *     and counted; duplicates collapse, the first occurrence fixes the position (PG:182);
 *     self-loops are legal;
 *   - entry ids: insertion order kept, duplicates and unknown ids ignored (PG:194-196);
 *   - parts: vertex ranges [part_off[p], part_off[p+1]) of the projects held in this graph
 *     (GraphService keeps one ProjectGraph per project, GS:34-38).  No edge may cross a
 *     range (DMG_EINVAL).  n_parts == 0 means one project.  A traversal query runs inside
 *     the project of its first valid source; sources in other projects match nothing
 *     (CCS:1426-1434, 1558).
 *   - names are optional (NULL name_off) for graphs that are only traversed.  lower_bytes is
 *     the host's own toLowerCase() of name_bytes with the SAME offsets (CCS:1513); if NULL
 *     the engine folds ASCII A-Z only.
 */
typedef struct dmg_graph_desc {
  uint32_t n_vertices;
  uint32_t n_entry;
  uint32_t n_parts;
  uint32_t reserved;
  uint64_t n_edges;
  const uint64_t* name_off;   /* [n_vertices+1] byte offsets into name_bytes, or NULL */
  const uint8_t* name_bytes;  /* UTF-8 */
  const uint8_t* lower_bytes; /* pre-lowered copy or NULL */
  const uint32_t* edge_src;   /* [n_edges], addDependency order */
  const uint32_t* edge_dst;
  const uint32_t* entry_ids;  /* [n_entry], markAsEntryPoint order */
  const uint32_t* part_off;   /* [n_parts+1] or NULL */
} dmg_graph_desc;

typedef struct dmg_graph_info {
  uint32_t n_vertices;
  uint32_t n_entry;
  uint32_t n_parts;
  uint32_t max_part_vertices;
  uint64_t n_edges;         /* kept */
  uint64_t dropped_edges;   /* unknown endpoint */
  uint64_t duplicate_edges; /* collapsed */
  uint64_t name_bytes;
  uint64_t device_bytes;
} dmg_graph_info;

/* Counters of the calling thread's most recent query call on this engine. */
typedef struct dmg_call_stats {
  uint64_t kernel_launches;  /* our kernels launched by the call */
  uint64_t edges_traversed;  /* TE of SURVEY.md 8d: sum of deg_mode(u) over every expanded frontier
                                vertex -- the work a top-down traversal does, whatever we did */
  uint64_t edges_read;       /* adjacency entries this call actually loaded (bottom-up levels
                                stop at the first frontier member, so this can be far smaller) */
  uint64_t rows_expanded;    /* vertices whose rows were read */
  uint64_t results;          /* entries written */
  uint64_t h2d_bytes;
  uint64_t d2h_bytes;
  float device_ms;           /* CUDA-event time of the device work on the call's stream */
  float reserved;
} dmg_call_stats;

/* ---- engine (GS lifecycle: the process-wide cache, GS:34-76) ------------------------------- */
int32_t dmg_device_count(void);
int32_t dmg_engine_create(const dmg_engine_opts* opts, dmg_engine** out);
void dmg_engine_destroy(dmg_engine* e);
/* Thread-local, never NULL; valid until the thread's next failing call. */
const char* dmg_last_error(void);
const char* dmg_version(void);

/* ---- graph handles (GS:84, 93-122: put / reload swap whole graphs) ------------------------- */
int32_t dmg_graph_create(dmg_engine* e, const dmg_graph_desc* desc, dmg_graph** out);
void dmg_graph_retain(dmg_graph* g);
void dmg_graph_release(dmg_graph* g); /* frees device memory when the count reaches 0 */
int32_t dmg_graph_info_get(const dmg_graph* g, dmg_graph_info* out);

/* Binary CSR snapshot: dmg_graph_save writes the graph exactly as the
 * device holds it (forward + reverse CSR, project ranges, entry list, names and their lower-cased
 * copy); dmg_graph_load uploads it again without JSON parsing, string hashing of edges or CSR
 * construction -- the startup path GraphService.loadAll (GS:55-76) pays per project today.  The
 * file is host-endian and meant as a sidecar of graph_data, rebuilt whenever the graph is.      */
int32_t dmg_graph_save(dmg_graph* g, const char* path);
int32_t dmg_graph_load(dmg_engine* e, const char* path, dmg_graph** out);

/* ---- T0 / T1: adjacency rows ---------------------------------------------------------------
 * dependencies(id) (PG:299-305): dir = DMG_OUT, row order = addDependency order.
 * dependents(id)   (PG:316-328): dir = DMG_IN, row order = ascending source id (the
 *   reference's order here is HashMap bucket order, pinned by no test).
 * Also serves vertexDeps (GQS:298-322), vertexOverview (GQS:243-265) and the all-vertex dump
 * of executeClasses (GQS:111-150) with ids = 0..V-1.
 * out_off has n+1 entries; row i is out_ids[out_off[i] .. out_off[i+1]).                       */
int32_t dmg_adjacency_batch(dmg_graph* g, int32_t dir, uint64_t n, const uint32_t* ids,
                            uint64_t* out_off, uint32_t* out_ids, uint64_t cap, uint64_t* needed);

/* ---- T2 / T3 / T4: batched multi-source k-hop neighbour resolution -------------------------
 * Query q has sources q_src[q_off[q] .. q_off[q+1]) (duplicates allowed; CCS:1563-1568 dedups
 * them through a HashSet).  Levels: L0 = distinct valid sources,
 * L(i+1) = (U_{u in Li} step(u)) minus everything seen so far, step = dependencies /
 * dependents / resolve for mode OUT / IN / BOTH.  Row q = U_{1<=i<=k} Li sorted by ascending
 * vertex id, out_hop = i.  exclude_sources != 0 seeds "seen" with L0, which at k = 1, mode
 * BOTH is exactly resolveNeighbors (CCS:1558-1578); exclude_sources == 0 with one source and
 * k = 1 is exactly resolve()/dependencies()/dependents() as sets (a self-loop reports the
 * vertex itself, PG test :829-837).  1 <= k <= DMG_MAX_HOPS.
 * Host pointers in, host pointers out.  How the rows travel is the library's business: a fetch call of
 * >= 2,048 queries on projects of <= 4096 vertices (k <= 7) crosses PCIe as bit rows -- 20x fewer bytes
 * than the packed rows -- and is expanded into out_ids / out_hop by the library's host threads; everything
 * else is packed on the device and copied.  The rows are the same either way.                      */
int32_t dmg_neighbors_batch(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources,
                            uint64_t n_queries, const uint32_t* q_off, const uint32_t* q_src,
                            uint64_t* out_off, uint32_t* out_ids, uint8_t* out_hop, uint64_t cap,
                            uint64_t* needed);

/* Same, every pointer a DEVICE pointer, nothing copied; returns after the work completed on
 * the device.  *needed (host) is always filled.  This is what bench.py times as `value`.       */
int32_t dmg_neighbors_batch_dev(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources,
                                uint64_t n_queries, const uint32_t* d_q_off, const uint32_t* d_q_src,
                                uint64_t* d_out_off, uint32_t* d_out_ids, uint8_t* d_out_hop,
                                uint64_t cap, uint64_t* needed);

/* Dense form of the same result, for graphs whose projects have <= 8192 vertices: row q is
 * *out_row_stride bytes, byte (v - lo_q) = hop of vertex v (0 = not a neighbour, 255 = excluded
 * source), with out_win[2q] = lo_q (first vertex of the query's project) and out_win[2q+1] = the
 * project's vertex count; bytes past the project's end are 0.  When k hops cover most of a project
 * (BASELINE config 2: ~1,990 of 2,000 vertices) this is 5x fewer bytes over PCIe than (id, hop)
 * pairs, and the host walks the row in id order.  cap_bytes / needed_bytes count bytes of out_rows. */
int32_t dmg_neighbors_batch_rows(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources,
                                 uint64_t n_queries, const uint32_t* q_off, const uint32_t* q_src,
                                 uint32_t* out_row_stride, uint32_t* out_win, uint8_t* out_rows,
                                 uint64_t cap_bytes, uint64_t* needed_bytes);

/* Compact form of the same result -- what a host should use when it walks the neighbours instead of
 * copying them (resolveNeighbors feeds every neighbour to a repository lookup, CCS:1583-1610): row q is
 * *out_planes bit planes of *out_plane_words uint32 words over the vertices of the query's project
 * (out_win[2q] = first vertex lo_q, out_win[2q+1] = vertex count); bit (v - lo_q) of plane b -- word
 * (v - lo_q) >> 5, bit (v - lo_q) & 31 -- is bit b of hop(v), so hop(v) = sum_b plane_b[v] << b and
 * 0 means "not a neighbour" (excluded sources included).  k <= 3 needs two planes: 512 bytes per query on
 * 2,000-vertex projects against ~10 KB of (id, hop) pairs -- 20x fewer PCIe bytes on BASELINE config 2.
 * out_counts[q] = number of neighbours of row q.  Needs projects of <= 4096 vertices and k <= 7
 * (DMG_EINVAL otherwise).  cap_bytes / needed_bytes count bytes of out_bits.  dmg_bits_decode turns such
 * rows into the packed rows of dmg_neighbors_batch on the host.                                      */
int32_t dmg_neighbors_batch_bits(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources,
                                 uint64_t n_queries, const uint32_t* q_off, const uint32_t* q_src,
                                 uint32_t* out_planes, uint32_t* out_plane_words, uint32_t* out_win,
                                 uint32_t* out_counts, uint32_t* out_bits, uint64_t cap_bytes,
                                 uint64_t* needed_bytes);
/* Host-side expansion of bit rows into CSR-shaped (id, hop) rows, ascending id (no device work; uses
 * n_threads host threads, 0 = all).  out_off has n_queries + 1 entries; cap counts entries.  Large results
 * are written with non-temporal stores from a per-thread staging buffer, and the bit rows that have been
 * consumed are flushed from the CPU caches (their contents are untouched): the buffer is normally the
 * target of the next dmg_neighbors_batch_bits copy, which is several times slower into cached lines.  */
int32_t dmg_bits_decode(uint32_t planes, uint32_t plane_words, uint64_t n_queries, const uint32_t* win,
                        const uint32_t* counts, const uint32_t* bits, uint64_t* out_off, uint32_t* out_ids,
                        uint8_t* out_hop, uint64_t cap, uint64_t* needed, int32_t n_threads);

/* The same k-hop batch for callers that must not allocate.  It has no counterpart in the
 * reference.  Some hosts cannot afford, or are not allowed, a cudaMalloc on the query path: a
 * request thread with a latency budget (device allocation can stall behind all other work on the
 * GPU), a process whose device memory is already carved up by another component's pool so that
 * a further allocation may fail under load, or code that runs where the CUDA runtime forbids
 * allocation.  This entry point calls no cudaMalloc / cudaFree and no allocating library
 * routine: all of its scratch (counters, scan status, row sizes, staging rows) is a 16 MiB
 * array compiled into the library, resident on every GPU that loads it until the process
 * exits.  The caller still passes device inputs and outputs exactly as for
 * dmg_neighbors_batch_dev, and the results are identical.
 *   - Limits: every project of the graph must have <= 8192 vertices, and
 *     n_queries * (round_up(largest project, 128) + 12) + ~1 KiB must fit in 16 MiB (about 8,100
 *     queries over 2,000-vertex projects); anything larger is rejected with DMG_EINVAL and a
 *     message that states both sizes.  Split the batch, or use the allocating entry point.
 *   - The first call on each device looks up the device address of the array; make that first
 *     call before starting any stream capture.
 *   - All calls share the one array, which keeps whatever the last call left there (the call
 *     clears the parts that must start at zero).  Two calls that overlap on the same GPU race
 *     on it: serialise them.  Work is issued on the default stream; the call returns after it
 *     completed.                                                                              */
int32_t dmg_neighbors_batch_dev_static(dmg_graph* g, int32_t mode, int32_t k, int32_t exclude_sources,
                                       uint64_t n_queries, const uint32_t* d_q_off, const uint32_t* d_q_src,
                                       uint64_t* d_out_off, uint32_t* d_out_ids, uint8_t* d_out_hop,
                                       uint64_t cap, uint64_t* needed);

/* ---- T5: closure and analysisOrder ----------------------------------------------------------
 * dmg_closure: BFS levels from a source set over dir (DMG_OUT / DMG_IN) on the whole graph;
 * out_level[v] = hop distance, 0 for sources, DMG_LEVEL_UNREACHED otherwise.  *reached counts
 * vertices with a level.  This is the reachable set of PG:390-421 with its BFS depth.
 * dmg_analysis_order: exactly ProjectGraph.analysisOrder() (PG:390-421): FIFO BFS from the
 * entry points in insertion order along dependencies() order, then every unvisited vertex in
 * nodes order.  out_order has n_vertices entries.                                              */
int32_t dmg_closure(dmg_graph* g, int32_t dir, uint32_t n_src, const uint32_t* src,
                    uint32_t* out_level, uint64_t* reached);
int32_t dmg_analysis_order(dmg_graph* g, uint32_t* out_order);

/* ---- vertex-range partitioned closure (BASELINE config 5) ---------------------
 * For the graph that is split over the GPUs of one box.  Rank r owns vertices [v_lo, v_hi)
 * (v_lo a multiple of 32) and creates its dmg_graph from the FULL vertex count but only the
 * edges with an endpoint in its range, so the rows of its own vertices are complete.  Frontier
 * and visited sets are V-bit bitmaps (uint32 words, bit v&31 of word v>>5) replicated on every
 * rank in DEVICE memory the host owns.  One level of the reference's BFS (PG:403-413) is:
 *   push  each rank ORs the `dir` neighbours of the frontier vertices it owns into its own
 *         full-width contribution bitmap (cleared by the call)            -> allgather all of them
 *   pull  each rank finds, for every unvisited vertex it owns, one frontier member in the
 *         vertex's reverse row and writes words [v_lo/32, ceil(v_hi/32)) of the next bitmap
 *         into d_next_slice                                               -> allgather the slices
 *   merge next = OR(gathered) & ~visited; visited |= next; the owner stamps `lvl` into
 *         d_level_slice[v - v_lo]; *n_new = |next| (identical on all ranks)
 * The exchange between expand and merge is the host's (ncclAllGather over NVLink in
 * dmgraph/dist.py); push suits small frontiers, pull large ones; the results are the same.    */
int32_t dmg_closure_push_dev(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi,
                             const uint32_t* d_frontier_bits, uint32_t* d_contrib_bits, uint64_t* edges);
int32_t dmg_closure_pull_dev(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi,
                             const uint32_t* d_frontier_bits, const uint32_t* d_visited_bits,
                             uint32_t* d_next_slice, uint64_t* edges);
int32_t dmg_closure_merge_dev(dmg_graph* g, uint32_t v_lo, uint32_t v_hi, uint32_t lvl,
                              const uint32_t* d_gathered, uint32_t n_gathered, uint32_t stride_words,
                              uint32_t* d_visited_bits, uint32_t* d_next_bits, uint32_t* d_level_slice,
                              uint64_t* n_new);

/* Stream-ordered forms of the three level primitives: work is only ENQUEUED on `cuda_stream` (a
 * cudaStream_t the host owns, e.g. the stream its NCCL calls run on); nothing is synchronised or
 * copied to the host.  *d_edges (device, 8 bytes) accumulates adjacency entries read; *d_n_new
 * (device, 8 bytes) receives |next| -- the host reads it once per level to test for termination.   */
int32_t dmg_closure_push_async(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi,
                               const uint32_t* d_frontier_bits, uint32_t* d_contrib_bits,
                               unsigned long long* d_edges, void* cuda_stream);
int32_t dmg_closure_pull_async(dmg_graph* g, int32_t dir, uint32_t v_lo, uint32_t v_hi,
                               const uint32_t* d_frontier_bits, const uint32_t* d_visited_bits,
                               uint32_t* d_next_slice, unsigned long long* d_edges, void* cuda_stream);
int32_t dmg_closure_merge_async(dmg_graph* g, uint32_t v_lo, uint32_t v_hi, uint32_t lvl,
                                const uint32_t* d_gathered, uint32_t n_gathered, uint32_t stride_words,
                                uint32_t* d_visited_bits, uint32_t* d_next_bits, uint32_t* d_level_slice,
                                unsigned long long* d_n_new, void* cuda_stream);

/* entryPoints() (PG:345-347), insertion order; isEntryPoint (PG:336-338) as a byte per id. */
int32_t dmg_entry_points(dmg_graph* g, uint32_t* out_ids, uint32_t cap, uint32_t* needed);
int32_t dmg_is_entry_point_batch(dmg_graph* g, uint64_t n, const uint32_t* ids, uint8_t* out_flags);

/* ---- S1 / S2: name scans --------------------------------------------------------------------
 * Queries are byte strings the host has already lower-cased (CCS:1507, GQS:431); query q is
 * q_bytes[q_off[q] .. q_off[q+1]).  The scan covers vertices [v_lo, v_hi) (one project's
 * range, or 0..V).  kind DMG_SEARCH_CONTAINS_ALL: row q = every id whose lower-cased
 * identifier contains the query, ascending id == nodes order (CCS:1511-1526; an empty query
 * matches everything).  The two *_FIRST kinds return at most one id per query (GQS:434-453).  */
int32_t dmg_search_batch(dmg_graph* g, int32_t kind, uint32_t v_lo, uint32_t v_hi, uint64_t n_queries,
                         const uint32_t* q_off, const uint8_t* q_bytes, uint64_t* out_off,
                         uint32_t* out_ids, uint64_t cap, uint64_t* needed);

/* ---- name -> id lookups on the device ---------------------------------
 * DMG_LOOKUP_EXACT: out_ids[i] = the vertex whose identifier equals text i byte for byte
 *   (nodes.containsKey / contains, PG:464; the per-frame class lookup that precedes
 *   resolveNeighbors, CCS:1404-1438, without one host HashMap probe or SQL round trip per frame);
 * DMG_LOOKUP_SIMPLE_FIRST: text i is a LOWER-CASED simple name; out_ids[i] = the first vertex in
 *   nodes order whose lower-cased simple name (text after the last '.') equals it (GQS:434-446).
 * Only vertices in [v_lo, v_hi) are considered (one project of a multi-project handle: the same
 * FQCN may exist in two projects).  0xFFFFFFFF = not found.                   */
#define DMG_LOOKUP_EXACT 0
#define DMG_LOOKUP_SIMPLE_FIRST 1
int32_t dmg_lookup_batch(dmg_graph* g, int32_t kind, uint32_t v_lo, uint32_t v_hi, uint64_t n,
                         const uint32_t* t_off, const uint8_t* t_bytes, uint32_t* out_ids);

/* ---- get_stack_trace_context, graph half, in one call --------------------
 * Replaces the per-frame lookups and the neighbour resolution of CodeContextService.getStackTraceContext
 * (CCS:1392-1446) and resolveNeighbors (CCS:1550-1578).  Trace t has frames t_off[t] .. t_off[t+1]; frame
 * f has class name cn_bytes[cn_off[f] .. cn_off[f+1]) and, if mn_off is given, method name
 * mn_bytes[mn_off[f] .. mn_off[f+1]) -- both as written (the reference matches them with SQL equality).
 *   out_status[f]  DMG_FRAME_MISSING        no project holds the class        (CCS:1405-1410)
 *                  DMG_FRAME_CLASS_MISSING  class known, method not           (CCS:1415-1423); only produced
 *                                           when mn_off is given AND the graph has dmg_graph_set_method_names
 *                  DMG_FRAME_FOUND          both found
 *   out_vertex[f]  the class's vertex (the lowest id if two projects hold the same name; the reference's
 *                  findOne() raises there, SURVEY quirk 11 -- a host concern), 0xFFFFFFFF if missing
 *   out_method[f]  catalog index of the method, 0xFFFFFFFF if none (may be NULL)
 *   row t          neighbours of the trace exactly as dmg_neighbors_batch(DMG_BOTH, k, exclude = 1) over the
 *                  FOUND frames: the graph is the project of the first found frame (CCS:1426-1434); a found
 *                  frame of another project takes part by name only (CCS:1563-1568).  k = 1 is the reference. */
#define DMG_FRAME_MISSING 0
#define DMG_FRAME_CLASS_MISSING 1
#define DMG_FRAME_FOUND 2
int32_t dmg_stacktrace_batch(dmg_graph* g, int32_t k, uint64_t n_traces, const uint32_t* t_off,
                             const uint32_t* cn_off, const uint8_t* cn_bytes, const uint32_t* mn_off,
                             const uint8_t* mn_bytes, uint8_t* out_status, uint32_t* out_vertex,
                             uint32_t* out_method, uint64_t* out_off, uint32_t* out_ids, uint8_t* out_hop,
                             uint64_t cap, uint64_t* needed);
/* Method names as written (source_methods.method_name), entry m = bytes[name_off[m] .. name_off[m+1]);
 * same catalog order as dmg_graph_set_methods, which must come first.                              */
int32_t dmg_graph_set_method_names(dmg_graph* g, const uint32_t* name_off, const uint8_t* bytes);

/* ---- Q1: graph_query execution (GQS:53-73, 111-150, 195-394) ---------------------------------
 * The DSL lexer (GraphQuery.parse, GQ:95-163) stays in the host; what reaches the engine is a
 * batch of already-tokenised operations.  Row i of the output answers ops[i]:
 *   DMG_OP_DEPENDENCIES  vertexDeps(outgoing)  GQS:298-322  ids in addDependency order, aux 0
 *   DMG_OP_DEPENDENTS    vertexDeps(incoming)  GQS:298-322  ids ascending, aux 0
 *   DMG_OP_OVERVIEW      vertexOverview        GQS:243-265  dependencies (aux 0) then dependents (aux 1)
 *   DMG_OP_KHOP          directed / undirected k-hop extension of the same primitives (SURVEY.md
 *                        8a T4; k = 1 is the reference): mode in `mode`, hops in `k`, aux = hop
 *   DMG_OP_ENTRYPOINTS   executeEntrypoints    GQS:152-191  entry ids in insertion order, aux 0
 *   DMG_OP_RESOLVE_NAME  resolveClassName steps 2 and 3, GQS:431-453 (the exact-id step 1 is the
 *                        host's map lookup): text = lower-cased target; at most one id; aux = 2 when
 *                        the simple name matched, 3 when only the containment scan did
 *   DMG_OP_METHOD        executeMethodCheck / vertexSingleMethod, GQS:324-394: text = lower-cased
 *                        method name; the row holds the catalog index of the first method of
 *                        `vertex` whose lower-cased name equals it (empty row = not found)
 *   DMG_OP_ENDPOINTS     executeEndpoints / ProjectGraph.allEndpoints, GQS:77-109, PG:606-618: catalog
 *                        indices of the methods flagged as HTTP endpoints (dmg_graph_set_method_flags),
 *                        ascending = vertex order then method order (the reference walks a HashMap
 *                        here; its order is pinned by no test)
 * text of an op is text_bytes[text_begin .. text_end).  Unknown vertices give empty rows.
 * On a handle that holds several projects, ENTRYPOINTS, ENDPOINTS and RESOLVE_NAME answer for the
 * project named by op.project only (the reference resolves the project first and then works on
 * that project's ProjectGraph, GQS:56-62, 423-456); vertex-addressed ops need no scope.         */
#define DMG_OP_DEPENDENCIES 0
#define DMG_OP_DEPENDENTS 1
#define DMG_OP_OVERVIEW 2
#define DMG_OP_KHOP 3
#define DMG_OP_ENTRYPOINTS 4
#define DMG_OP_RESOLVE_NAME 5
#define DMG_OP_METHOD 6
#define DMG_OP_ENDPOINTS 7

typedef struct dmg_query_op {
  uint32_t opcode;
  uint32_t vertex;
  uint32_t mode;            /* DMG_OP_KHOP: DMG_OUT / DMG_IN / DMG_BOTH */
  uint32_t k;               /* DMG_OP_KHOP: hops */
  uint32_t exclude_sources; /* DMG_OP_KHOP */
  uint32_t text_begin;
  uint32_t text_end;
  uint32_t project;         /* DMG_OP_ENTRYPOINTS / ENDPOINTS / RESOLVE_NAME: index of the project range
                               (dmg_graph_desc.part_off) the op is asked of -- graph_query names ONE
                               project (GQS:56-62); 0 on a single-project handle; out of range = DMG_EINVAL */
} dmg_query_op;

int32_t dmg_query_batch(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                        uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint64_t cap, uint64_t* needed);

/* Decorated rows: the same batch with one more byte per returned entry, so that
 * vertexDeps / executeClasses rows (GQS:122-130, 308-316) need no nodeInfo / entry-point lookup per result
 * on the host: out_meta = class-type code (bits 0-6, as given to dmg_graph_set_class_types; 0 = no
 * nodeInfo) | 0x80 if the vertex is an entry point (PG:336-338).  Entries that are method-catalog indices
 * (DMG_OP_METHOD, DMG_OP_ENDPOINTS) get 0.  description / sourceFile stay host strings indexed by the id. */
int32_t dmg_query_batch_meta(dmg_graph* g, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                             uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint8_t* out_meta,
                             uint64_t cap, uint64_t* needed);
/* One code per vertex (NodeInfo.classType, PG:504-510, as a host-defined enumeration 1..127; 0 = none). */
int32_t dmg_graph_set_class_types(dmg_graph* g, const uint8_t* codes);

/* Method catalog of a graph (ProjectGraph.methodInfoMap, PG:112; names only -- descriptions and
 * business logic stay host payload): methods of vertex v are catalog entries
 * [method_off[v], method_off[v+1]), entry m's lower-cased name is
 * lower_bytes[name_off[m] .. name_off[m+1]).  Optional; needed only by DMG_OP_METHOD.           */
int32_t dmg_graph_set_methods(dmg_graph* g, const uint32_t* method_off, const uint32_t* name_off,
                              const uint8_t* lower_bytes);
/* One byte per catalog entry; bit 0 = MethodInfo.isHttpEndpoint() (httpMethod and httpPath present,
 * PG:75-77).  Needs dmg_graph_set_methods first.                                                 */
int32_t dmg_graph_set_method_flags(dmg_graph* g, const uint8_t* flags);

/* ---- one process, N devices ------------------------------------------------------
 * The reference is one JVM with one registry of graphs (GS:34-38): multi-GPU has to live behind this ABI,
 * not in a launcher.  dmg_dist_create makes one engine per listed device (devices == NULL: 0 .. n-1).
 *   replicated graph   (partitioned = 0) every device holds the whole graph; dmg_dist_neighbors_batch cuts
 *                      the batch into n contiguous query ranges, one host thread per device, rows come back
 *                      in query order exactly as from dmg_neighbors_batch.  No collective.
 *   partitioned graph  (partitioned = 1) rank r owns a vertex range (multiples of 32) and the edges with
 *                      an endpoint in it; dmg_dist_closure runs the BFS of analysisOrder (PG:390-421) level
 *                      by level.  Transport 1 (devices with peer access to each other, the NVSwitch case):
 *                      one persistent kernel per device for the whole closure; sparse levels send one
 *                      4-byte remote reduction per newly met vertex into the owner's slice, owners store
 *                      the changed frontier / visited words into every replica, two flag barriers per level
 *                      in peer memory, no host round trip.  Transport 0 (no peer access, or DMG_DIST_NCCL=1
 *                      in the environment): an ncclAllGather of frontier bitmaps per level (communicators
 *                      are created in-process on first use; NCCL is bound with dlopen).  out_level[v] = hop
 *                      distance, DMG_LEVEL_UNREACHED if not reached -- identical to dmg_closure.
 * NCCL missing or failing: DMG_ENCCL with the NCCL error text in dmg_last_error().  A device that does not
 * reach a level barrier of transport 1 within 2 s ends the call with DMG_ECUDA instead of hanging.       */
typedef struct dmg_dist dmg_dist;
typedef struct dmg_dist_graph dmg_dist_graph;
typedef struct dmg_dist_closure_stats {
  uint32_t levels, push_levels, pull_levels;
  uint32_t transport;                 /* 0 = NCCL allgather per level, 1 = persistent kernel over peer memory */
  uint64_t edges_read;                /* adjacency entries loaded, all ranks */
  uint64_t exchanged_bytes_per_rank;  /* transport 0: received through the allgathers; 1: largest sender's bytes */
  float device_ms;                    /* CUDA events on rank 0's stream around the level loop / its kernel */
  float reserved2;
} dmg_dist_closure_stats;
int32_t dmg_dist_create(const int32_t* devices, int32_t n, dmg_dist** out);
void dmg_dist_destroy(dmg_dist* d);
int32_t dmg_dist_size(const dmg_dist* d);
int32_t dmg_dist_engine(dmg_dist* d, int32_t rank, dmg_engine** out);
int32_t dmg_dist_graph_create(dmg_dist* d, const dmg_graph_desc* desc, int32_t partitioned, dmg_dist_graph** out);
void dmg_dist_graph_release(dmg_dist_graph* dg);
int32_t dmg_dist_graph_part(const dmg_dist_graph* dg, int32_t rank, dmg_graph** out, uint32_t* v_lo, uint32_t* v_hi);
int32_t dmg_dist_neighbors_batch(dmg_dist_graph* dg, int32_t mode, int32_t k, int32_t exclude_sources,
                                 uint64_t n_queries, const uint32_t* q_off, const uint32_t* q_src,
                                 uint64_t* out_off, uint32_t* out_ids, uint8_t* out_hop, uint64_t cap,
                                 uint64_t* needed);
/* dmg_query_batch (GQS:53-73, 111-150, 195-394) over the replicas of a replicated graph: contiguous op
 * ranges, one host thread per device, rows and offsets exactly as one dmg_query_batch call gives them.
 * Method catalogs / flags / class types must have been set on every replica (dmg_dist_graph_part).      */
int32_t dmg_dist_query_batch(dmg_dist_graph* dg, uint64_t n_ops, const dmg_query_op* ops, const uint8_t* text_bytes,
                             uint64_t* out_off, uint32_t* out_ids, uint8_t* out_aux, uint64_t cap, uint64_t* needed);
int32_t dmg_dist_closure(dmg_dist_graph* dg, int32_t dir, uint32_t n_src, const uint32_t* src,
                         uint32_t* out_level, uint64_t* reached, dmg_dist_closure_stats* stats);
/* analysisOrder() (PG:390-421) of a PARTITIONED graph, exactly as dmg_analysis_order gives it for one device:
 * entry points in insertion order, then FIFO discovery order, then the unreached vertices in nodes order
 * (PG:415-418).  Runs dmg_dist_closure for the depths, then orders level by level: the owner of a parent
 * pushes (parent's place in its level, edge position) as a 64-bit minimum into the child's owner, counts and
 * places its winning edges through peer memory.  out_order has n_vertices entries.  Needs peer access
 * between all devices (DMG_EINVAL otherwise).                                                              */
int32_t dmg_dist_analysis_order(dmg_dist_graph* dg, uint32_t* out_order);

/* ---- measurement -----------------------------------------------------------------------------*/
int32_t dmg_last_call_stats(dmg_engine* e, dmg_call_stats* out);
/* Per-kernel CUDA-event timing (events on the stream each kernel is launched on).  While
 * enabled every launch is bracketed by two events; read the totals with dmg_profile_get.    */
int32_t dmg_profile_enable(dmg_engine* e, int32_t on);
int32_t dmg_profile_reset(dmg_engine* e);
int32_t dmg_profile_count(dmg_engine* e);
int32_t dmg_profile_get(dmg_engine* e, int32_t index, const char** kernel_name, uint64_t* launches,
                        double* total_ms);

/* ---- device / pinned memory helpers (so a host needs no CUDA binding of its own) ------------ */
int32_t dmg_device_alloc(dmg_engine* e, uint64_t bytes, void** d_ptr);
int32_t dmg_device_free(dmg_engine* e, void* d_ptr);
int32_t dmg_copy_to_device(dmg_engine* e, void* d_dst, const void* src, uint64_t bytes);
int32_t dmg_copy_to_host(dmg_engine* e, void* dst, const void* d_src, uint64_t bytes);
int32_t dmg_pinned_alloc(dmg_engine* e, uint64_t bytes, void** ptr);
int32_t dmg_pinned_free(dmg_engine* e, void* ptr);
/* Writes `bytes` of device memory (> L2) so the next timed call starts with a cold L2. */
int32_t dmg_flush_l2(dmg_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* DMGRAPH_H */
