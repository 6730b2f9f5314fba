/*
 * dmgraph_host.h -- host-side mirror of the reference's graph-facing classes, above the device
 * C ABI of dmgraph.h.  The reference is compiled (Java) code, so this layer is C++ inside
 * libdmgraph.so with a C surface; the Java shim may use it or keep its own
 * String<->id map and talk to dmgraph.h directly.
 *
 *   dmg_pgraph   ~ ProjectGraph under construction: nodes / edges / entryPoints exactly as
 *                  addNode / addDependency / markAsEntryPoint leave them (PG:160-197), plus the
 *                  wire format toJson / fromJson (PG:634-905).
 *   dmg_service  ~ GraphService (GS:34-146): projectId -> published device graph, name index,
 *                  put / reload swap whole graphs; readers retain the handle they get.
 *
 * Strings are NUL-terminated UTF-8.  "blank" follows String.isBlank() for the ASCII range
 * (shared/Preconditions.java:41-47) and yields DMG_EINVAL where Java throws
 * IllegalArgumentException.
 */
#ifndef DMGRAPH_HOST_H
#define DMGRAPH_HOST_H

#include "dmgraph.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dmg_pgraph dmg_pgraph;
typedef struct dmg_service dmg_service;

/* ---- ProjectGraph (host half) -------------------------------------------------------------- */
dmg_pgraph* dmg_pg_new(void);                                      /* PG:117-125 */
void dmg_pg_free(dmg_pgraph* pg);
int32_t dmg_pg_add_node(dmg_pgraph* pg, const char* id, const char* source_file);  /* PG:160-164 */
int32_t dmg_pg_add_dependency(dmg_pgraph* pg, const char* from, const char* to);   /* PG:175-183 */
int32_t dmg_pg_mark_entry_point(dmg_pgraph* pg, const char* id);                   /* PG:190-197 */
uint32_t dmg_pg_node_count(const dmg_pgraph* pg);                                  /* PG:473 */
uint32_t dmg_pg_entry_point_count(const dmg_pgraph* pg);                           /* PG:482 */
uint64_t dmg_pg_edge_count(const dmg_pgraph* pg);
int32_t dmg_pg_contains(const dmg_pgraph* pg, const char* id);                     /* PG:464 */
/* vertex id of an identifier, or -1 (the id contract: position in `nodes`, PG:118) */
int64_t dmg_pg_id_of(const dmg_pgraph* pg, const char* id);
const char* dmg_pg_name_of(const dmg_pgraph* pg, uint32_t vertex);                 /* identifiers(), PG:491 */
const char* dmg_pg_source_file(const dmg_pgraph* pg, const char* id);              /* PG:454 */
/* Upload as one immutable device graph (names + pre-lowered ASCII copy included). */
int32_t dmg_pg_publish(dmg_engine* e, const dmg_pgraph* pg, dmg_graph** out);
/* Wire format (PG:634-750 / 758-905).  The three metadata sections (methodParameters, nodeInfo,
 * methodInfo) and per-node classId are carried through verbatim; old-format JSON without them
 * loads (ProjectGraphTest.java:982-992).  fromJson does not re-validate edges against nodes
 * (PG:782-795); edges with an unknown endpoint are kept on the host copy and dropped + counted
 * at publish.  *out_json is malloc'ed; free it with dmg_free_string. */
int32_t dmg_pg_to_json(const dmg_pgraph* pg, char** out_json);
int32_t dmg_pg_from_json(const char* json, dmg_pgraph** out);
void dmg_free_string(char* s);

/* ---- GraphService -------------------------------------------------------------------------- */
int32_t dmg_service_new(dmg_engine* e, dmg_service** out);
void dmg_service_free(dmg_service* s);
/* put (GS:118-122): publish `pg` and swap it in under project_id / project_name. */
int32_t dmg_service_put(dmg_service* s, const char* project_id, const char* project_name, const dmg_pgraph* pg);
/* reload / loadAll element (GS:55-76, 93-109): parse graph_data JSON, publish, swap in.  A graph
 * that fails to parse is skipped and left absent, as the reference logs and continues. */
int32_t dmg_service_load_json(dmg_service* s, const char* project_id, const char* project_name, const char* json);
/* getGraph (GS:84) / getGraphByProjectName (GS:130-136): retained handle or NULL.  The caller
 * releases it with dmg_graph_release; a concurrent put never frees it under the reader. */
dmg_graph* dmg_service_get(dmg_service* s, const char* project_id);
dmg_graph* dmg_service_get_by_name(dmg_service* s, const char* project_name);
uint32_t dmg_service_size(dmg_service* s);

#ifdef __cplusplus
}
#endif
#endif
