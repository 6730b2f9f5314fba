// Host-side mirror of ProjectGraph (build half + wire format) and GraphService, above the device
// C ABI.  No CUDA here: everything that computes a graph result goes through dmgraph.h.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/dmgraph_host.h"

namespace dmg {
void set_error(const char* fmt, ...);
}
using dmg::set_error;

namespace {

inline bool java_ws(unsigned char c) { return (c >= 9 && c <= 13) || (c >= 28 && c <= 32); }
bool is_blank(const char* s) {   // String.isBlank(), ASCII range (shared/Preconditions.java:41-47)
  if (!s) return true;
  for (; *s; ++s)
    if (!java_ws((unsigned char)*s)) return false;
  return true;
}

// ---- minimal JSON reader / writer for the graph_data document ----------------------------------
struct JsonIn {
  const char* p;
  const char* end;
  bool ok = true;
  void ws() {
    while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p;
  }
  bool eat(char c) {
    ws();
    if (p < end && *p == c) {
      ++p;
      return true;
    }
    return false;
  }
  static void put_utf8(std::string& o, uint32_t cp) {
    if (cp < 0x80) o += (char)cp;
    else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 63)); }
    else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 63)); o += (char)(0x80 | (cp & 63)); }
    else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 63)); o += (char)(0x80 | ((cp >> 6) & 63)); o += (char)(0x80 | (cp & 63)); }
  }
  bool hex4(uint32_t* v) {
    if (end - p < 4) return false;
    uint32_t x = 0;
    for (int i = 0; i < 4; ++i) {
      char c = p[i];
      x <<= 4;
      if (c >= '0' && c <= '9') x |= c - '0';
      else if (c >= 'a' && c <= 'f') x |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') x |= c - 'A' + 10;
      else return false;
    }
    p += 4;
    *v = x;
    return true;
  }
  bool string(std::string* out) {
    ws();
    if (p >= end || *p != '"') return ok = false;
    ++p;
    out->clear();
    while (p < end && *p != '"') {
      if (*p == '\\') {
        if (++p >= end) return ok = false;
        char c = *p++;
        switch (c) {
          case 'n': *out += '\n'; break;
          case 't': *out += '\t'; break;
          case 'r': *out += '\r'; break;
          case 'b': *out += '\b'; break;
          case 'f': *out += '\f'; break;
          case 'u': {
            uint32_t cp;
            if (!hex4(&cp)) return ok = false;
            if (cp >= 0xD800 && cp < 0xDC00 && end - p >= 6 && p[0] == '\\' && p[1] == 'u') {
              p += 2;
              uint32_t lo;
              if (!hex4(&lo)) return ok = false;
              cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            }
            put_utf8(*out, cp);
            break;
          }
          case '"': case '\\': case '/': *out += c; break;
          default: return ok = false;   // Jackson: "Unrecognized character escape"
        }
      } else {
        if ((unsigned char)*p < 0x20) return ok = false;   // Jackson: "Illegal unquoted character"
        *out += *p++;
      }
    }
    if (p >= end) return ok = false;
    ++p;
    return true;
  }
  // skips any value; returns its raw text
  bool skip(std::string* raw) {
    ws();
    const char* s = p;
    if (p >= end) return ok = false;
    if (*p == '"') {
      std::string t;
      if (!string(&t)) return false;
    } else if (*p == '{' || *p == '[') {
      const char open = *p, close = (open == '{') ? '}' : ']';
      ++p;
      ws();
      if (p < end && *p == close) {
        ++p;
      } else {
        while (true) {
          if (open == '{') {
            std::string k;
            if (!string(&k) || !eat(':')) return ok = false;
          }
          if (!skip(nullptr)) return false;
          if (eat(',')) continue;
          if (eat(close)) break;
          return ok = false;
        }
      }
    } else if (!scalar()) {
      return ok = false;
    }
    if (raw) raw->assign(s, p - s);
    return true;
  }
  // true / false / null / a JSON number, nothing else (Jackson rejects everything else too)
  bool scalar() {
    for (const char* lit : {"true", "false", "null"}) {
      const size_t n = strlen(lit);
      if ((size_t)(end - p) >= n && strncmp(p, lit, n) == 0) {
        p += n;
        return true;
      }
    }
    const char* q = p;
    if (q < end && *q == '-') ++q;
    if (q >= end || *q < '0' || *q > '9') return false;
    if (*q == '0') ++q;
    else while (q < end && *q >= '0' && *q <= '9') ++q;
    if (q < end && *q == '.') {
      ++q;
      if (q >= end || *q < '0' || *q > '9') return false;
      while (q < end && *q >= '0' && *q <= '9') ++q;
    }
    if (q < end && (*q == 'e' || *q == 'E')) {
      ++q;
      if (q < end && (*q == '+' || *q == '-')) ++q;
      if (q >= end || *q < '0' || *q > '9') return false;
      while (q < end && *q >= '0' && *q <= '9') ++q;
    }
    p = q;
    return true;
  }
};

void json_escape(std::string& o, const std::string& s) {
  o += '"';
  for (unsigned char c : s) {
    switch (c) {
      case '"': o += "\\\""; break;
      case '\\': o += "\\\\"; break;
      case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break;
      case '\t': o += "\\t"; break;
      case '\b': o += "\\b"; break;
      case '\f': o += "\\f"; break;
      default:
        if (c < 0x20) {
          char buf[8];
          snprintf(buf, sizeof(buf), "\\u%04X", c);
          o += buf;
        } else {
          o += (char)c;
        }
    }
  }
  o += '"';
}

}  // namespace

// One ProjectGraph as the host builds it.  `edges` keeps, per source, the targets in
// first-insertion order (LinkedHashSet, PG:182); sources are kept in first-insertion order too
// (the reference's HashMap order is unpinned).
struct dmg_pgraph {
  std::vector<std::string> names, files, class_ids;
  std::vector<char> has_class_id;   // classIdMapping.containsKey: an empty string is a value too (PG:641-644, 775-778)
  std::unordered_map<std::string, uint32_t> pos;
  struct Row {
    std::string src;
    std::vector<std::string> dst;
    std::unordered_set<std::string> has;
  };
  std::vector<Row> rows;
  std::unordered_map<std::string, uint32_t> row_of;
  std::vector<std::string> entry;
  std::unordered_set<std::string> entry_has;
  std::string raw_method_params = "{}", raw_node_info = "{}", raw_method_info = "{}";
  uint64_t n_edges = 0;

  void put_edge(const std::string& a, const std::string& b) {
    auto it = row_of.find(a);
    if (it == row_of.end()) {
      it = row_of.emplace(a, (uint32_t)rows.size()).first;
      rows.emplace_back();
      rows.back().src = a;
    }
    Row& r = rows[it->second];
    if (r.has.insert(b).second) {
      r.dst.push_back(b);
      ++n_edges;
    }
  }
};

struct dmg_service {
  dmg_engine* eng;
  std::mutex mu;
  std::unordered_map<std::string, dmg_graph*> graphs;       // projectId -> graph (one ref held)
  std::unordered_map<std::string, std::string> name_to_id;  // GS:37-38
};

extern "C" {

dmg_pgraph* dmg_pg_new(void) { return new (std::nothrow) dmg_pgraph(); }
void dmg_pg_free(dmg_pgraph* pg) { delete pg; }

int32_t dmg_pg_add_node(dmg_pgraph* pg, const char* id, const char* file) {
  if (!pg) return DMG_EINVAL;
  if (is_blank(id)) {
    set_error("Identifier is required");
    return DMG_EINVAL;
  }
  if (is_blank(file)) {
    set_error("Source file is required");
    return DMG_EINVAL;
  }
  auto it = pg->pos.find(id);
  if (it == pg->pos.end()) {
    pg->pos.emplace(id, (uint32_t)pg->names.size());
    pg->names.emplace_back(id);
    pg->files.emplace_back(file);
    pg->class_ids.emplace_back();
    pg->has_class_id.push_back(0);
  } else {
    pg->files[it->second] = file;   // LinkedHashMap.put: position kept, value replaced
  }
  return DMG_OK;
}

int32_t dmg_pg_add_dependency(dmg_pgraph* pg, const char* from, const char* to) {
  if (!pg) return DMG_EINVAL;
  if (is_blank(from)) {
    set_error("From identifier is required");
    return DMG_EINVAL;
  }
  if (is_blank(to)) {
    set_error("To identifier is required");
    return DMG_EINVAL;
  }
  if (!pg->pos.count(from) || !pg->pos.count(to)) return DMG_OK;   // silently ignored (PG:179-181)
  pg->put_edge(from, to);
  return DMG_OK;
}

int32_t dmg_pg_mark_entry_point(dmg_pgraph* pg, const char* id) {
  if (!pg) return DMG_EINVAL;
  if (is_blank(id)) {
    set_error("Entry point identifier is required");
    return DMG_EINVAL;
  }
  if (pg->pos.count(id) && pg->entry_has.insert(id).second) pg->entry.emplace_back(id);
  return DMG_OK;
}

uint32_t dmg_pg_node_count(const dmg_pgraph* pg) { return pg ? (uint32_t)pg->names.size() : 0; }
uint32_t dmg_pg_entry_point_count(const dmg_pgraph* pg) { return pg ? (uint32_t)pg->entry.size() : 0; }
uint64_t dmg_pg_edge_count(const dmg_pgraph* pg) { return pg ? pg->n_edges : 0; }
int32_t dmg_pg_contains(const dmg_pgraph* pg, const char* id) { return pg && id && pg->pos.count(id); }
int64_t dmg_pg_id_of(const dmg_pgraph* pg, const char* id) {
  if (!pg || !id) return -1;
  auto it = pg->pos.find(id);
  return it == pg->pos.end() ? -1 : (int64_t)it->second;
}
const char* dmg_pg_name_of(const dmg_pgraph* pg, uint32_t v) {
  return (pg && v < pg->names.size()) ? pg->names[v].c_str() : nullptr;
}
const char* dmg_pg_source_file(const dmg_pgraph* pg, const char* id) {
  const int64_t v = dmg_pg_id_of(pg, id);
  return v < 0 ? nullptr : pg->files[(size_t)v].c_str();
}

int32_t dmg_pg_publish(dmg_engine* e, const dmg_pgraph* pg, dmg_graph** out) {
  if (!e || !pg || !out) {
    set_error("dmg_pg_publish: NULL argument");
    return DMG_EINVAL;
  }
  const uint32_t V = (uint32_t)pg->names.size();
  std::vector<uint64_t> off(V + 1, 0);
  std::string arena;
  for (uint32_t v = 0; v < V; ++v) {
    arena += pg->names[v];
    off[v + 1] = arena.size();
  }
  std::vector<uint32_t> src, dst, ent;
  src.reserve(pg->n_edges);
  dst.reserve(pg->n_edges);
  auto id_or_unknown = [&](const std::string& s) {
    auto it = pg->pos.find(s);
    return it == pg->pos.end() ? 0xFFFFFFFFu : it->second;   // dangling -> dropped by the engine
  };
  for (const auto& r : pg->rows) {
    const uint32_t s = id_or_unknown(r.src);
    for (const auto& t : r.dst) {
      src.push_back(s);
      dst.push_back(id_or_unknown(t));
    }
  }
  for (const auto& x : pg->entry) ent.push_back(id_or_unknown(x));
  dmg_graph_desc d;
  memset(&d, 0, sizeof(d));
  d.n_vertices = V;
  d.n_edges = src.size();
  d.edge_src = src.data();
  d.edge_dst = dst.data();
  d.n_entry = (uint32_t)ent.size();
  d.entry_ids = ent.data();
  d.name_off = off.data();
  d.name_bytes = reinterpret_cast<const uint8_t*>(arena.data());
  d.lower_bytes = nullptr;   // ASCII fold by the engine; a Java host passes its own toLowerCase()
  return dmg_graph_create(e, &d, out);
}

int32_t dmg_pg_to_json(const dmg_pgraph* pg, char** out_json) {
  if (!pg || !out_json) return DMG_EINVAL;
  std::string o = "{\"nodes\":{";
  for (size_t v = 0; v < pg->names.size(); ++v) {
    if (v) o += ',';
    json_escape(o, pg->names[v]);
    o += ":{\"sourceFile\":";
    json_escape(o, pg->files[v]);
    if (pg->has_class_id[v]) {
      o += ",\"classId\":";
      json_escape(o, pg->class_ids[v]);
    }
    o += '}';
  }
  o += "},\"edges\":{";
  bool first = true;
  for (const auto& r : pg->rows) {
    if (!first) o += ',';
    first = false;
    json_escape(o, r.src);
    o += ":[";
    for (size_t i = 0; i < r.dst.size(); ++i) {
      if (i) o += ',';
      json_escape(o, r.dst[i]);
    }
    o += ']';
  }
  o += "},\"entryPoints\":[";
  for (size_t i = 0; i < pg->entry.size(); ++i) {
    if (i) o += ',';
    json_escape(o, pg->entry[i]);
  }
  o += "],\"methodParameters\":" + pg->raw_method_params + ",\"nodeInfo\":" + pg->raw_node_info +
       ",\"methodInfo\":" + pg->raw_method_info + "}";
  char* buf = (char*)malloc(o.size() + 1);
  if (!buf) return DMG_ENOMEM;
  memcpy(buf, o.c_str(), o.size() + 1);
  *out_json = buf;
  return DMG_OK;
}

int32_t dmg_pg_from_json(const char* json, dmg_pgraph** out) {
  if (!out) return DMG_EINVAL;
  *out = nullptr;
  if (is_blank(json)) {
    set_error("JSON is required");   // PG:759
    return DMG_EINVAL;
  }
  std::unique_ptr<dmg_pgraph> pg(new dmg_pgraph());
  JsonIn in{json, json + strlen(json)};
  if (!in.eat('{')) {
    set_error("Failed to deserialize project graph: not a JSON object");
    return DMG_EINVAL;
  }
  if (!in.eat('}')) {
    while (true) {
      std::string key;
      if (!in.string(&key) || !in.eat(':')) { in.ok = false; break; }
      if (key == "nodes") {
        if (!in.eat('{')) { in.ok = false; break; }
        if (!in.eat('}'))
          while (true) {
            std::string id, file, cid;
            bool have_file = false, have_cid = false;
            if (!in.string(&id) || !in.eat(':') || !in.eat('{')) { in.ok = false; break; }
            if (!in.eat('}'))
              while (true) {
                std::string k2, v2;
                if (!in.string(&k2) || !in.eat(':')) { in.ok = false; break; }
                in.ws();
                if (in.p < in.end && *in.p == '"') {
                  if (!in.string(&v2)) break;
                  if (k2 == "sourceFile") { file = v2; have_file = true; }
                  else if (k2 == "classId") { cid = v2; have_cid = true; }
                } else if (!in.skip(nullptr)) break;
                if (in.eat(',')) continue;
                if (in.eat('}')) break;
                in.ok = false;
                break;
              }
            if (!in.ok) break;
            if (!have_file) { in.ok = false; break; }   // value.get("sourceFile") would NPE
            auto it = pg->pos.find(id);
            if (it == pg->pos.end()) {
              pg->pos.emplace(id, (uint32_t)pg->names.size());
              pg->names.push_back(id);
              pg->files.push_back(file);
              pg->class_ids.push_back(cid);
              pg->has_class_id.push_back(have_cid ? 1 : 0);
            } else {
              pg->files[it->second] = file;
              pg->class_ids[it->second] = cid;
              pg->has_class_id[it->second] = have_cid ? 1 : 0;
            }
            if (in.eat(',')) continue;
            if (in.eat('}')) break;
            in.ok = false;
            break;
          }
      } else if (key == "edges") {
        if (!in.eat('{')) { in.ok = false; break; }
        if (!in.eat('}'))
          while (true) {
            std::string s;
            if (!in.string(&s) || !in.eat(':') || !in.eat('[')) { in.ok = false; break; }
            // edges.put(key, deps): a repeated key replaces the earlier set (PG:794)
            auto it = pg->row_of.find(s);
            if (it != pg->row_of.end()) {
              pg->n_edges -= pg->rows[it->second].dst.size();
              pg->rows[it->second].dst.clear();
              pg->rows[it->second].has.clear();
            } else {
              pg->row_of.emplace(s, (uint32_t)pg->rows.size());
              pg->rows.emplace_back();
              pg->rows.back().src = s;
            }
            if (!in.eat(']'))
              while (true) {
                std::string t;
                if (!in.string(&t)) break;
                pg->put_edge(s, t);   // not validated against nodes (PG:782-795)
                if (in.eat(',')) continue;
                if (in.eat(']')) break;
                in.ok = false;
                break;
              }
            if (!in.ok) break;
            if (in.eat(',')) continue;
            if (in.eat('}')) break;
            in.ok = false;
            break;
          }
      } else if (key == "entryPoints") {
        if (!in.eat('[')) { in.ok = false; break; }
        if (!in.eat(']'))
          while (true) {
            std::string t;
            if (!in.string(&t)) break;
            if (pg->entry_has.insert(t).second) pg->entry.push_back(t);
            if (in.eat(',')) continue;
            if (in.eat(']')) break;
            in.ok = false;
            break;
          }
      } else {
        std::string raw;
        if (!in.skip(&raw)) break;
        if (key == "methodParameters") pg->raw_method_params = raw;
        else if (key == "nodeInfo") pg->raw_node_info = raw;
        else if (key == "methodInfo") pg->raw_method_info = raw;
      }
      if (!in.ok) break;
      if (in.eat(',')) continue;
      if (in.eat('}')) break;
      in.ok = false;
      break;
    }
  }
  if (!in.ok) {
    set_error("Failed to deserialize project graph: malformed JSON near byte %ld", (long)(in.p - json));
    return DMG_EINVAL;
  }
  *out = pg.release();
  return DMG_OK;
}

void dmg_free_string(char* s) { free(s); }

// ---- GraphService --------------------------------------------------------------------------------
int32_t dmg_service_new(dmg_engine* e, dmg_service** out) {
  if (!e || !out) return DMG_EINVAL;
  *out = new (std::nothrow) dmg_service();
  if (!*out) return DMG_ENOMEM;
  (*out)->eng = e;
  return DMG_OK;
}

void dmg_service_free(dmg_service* s) {
  if (!s) return;
  for (auto& kv : s->graphs) dmg_graph_release(kv.second);
  delete s;
}

static int32_t service_swap(dmg_service* s, const char* id, const char* name, dmg_graph* g) {
  dmg_graph* old = nullptr;
  {
    std::lock_guard<std::mutex> lk(s->mu);
    auto it = s->graphs.find(id);
    if (it != s->graphs.end()) {
      old = it->second;
      it->second = g;
    } else {
      s->graphs.emplace(id, g);
    }
    if (name) s->name_to_id[name] = id;   // GS:66, 120
  }
  if (old) dmg_graph_release(old);   // readers that retained it keep it alive
  return DMG_OK;
}

int32_t dmg_service_put(dmg_service* s, const char* project_id, const char* project_name, const dmg_pgraph* pg) {
  if (!s || !project_id || !pg) {
    set_error("dmg_service_put: NULL argument");
    return DMG_EINVAL;
  }
  dmg_graph* g = nullptr;
  int32_t st = dmg_pg_publish(s->eng, pg, &g);
  if (st != DMG_OK) return st;
  return service_swap(s, project_id, project_name, g);
}

int32_t dmg_service_load_json(dmg_service* s, const char* project_id, const char* project_name, const char* json) {
  if (!s || !project_id) {
    set_error("dmg_service_load_json: NULL argument");
    return DMG_EINVAL;
  }
  dmg_pgraph* pg = nullptr;
  int32_t st = dmg_pg_from_json(json, &pg);
  if (st != DMG_OK) return st;   // GS:62-72: the caller logs and skips this project
  st = dmg_service_put(s, project_id, project_name, pg);
  dmg_pg_free(pg);
  return st;
}

dmg_graph* dmg_service_get(dmg_service* s, const char* project_id) {
  if (!s || !project_id) return nullptr;
  std::lock_guard<std::mutex> lk(s->mu);
  auto it = s->graphs.find(project_id);
  if (it == s->graphs.end()) return nullptr;
  dmg_graph_retain(it->second);
  return it->second;
}

dmg_graph* dmg_service_get_by_name(dmg_service* s, const char* project_name) {
  if (!s || !project_name) return nullptr;
  std::string id;
  {
    std::lock_guard<std::mutex> lk(s->mu);
    auto it = s->name_to_id.find(project_name);
    if (it == s->name_to_id.end()) return nullptr;   // GS:131-134
    id = it->second;
  }
  return dmg_service_get(s, id.c_str());
}

uint32_t dmg_service_size(dmg_service* s) {
  if (!s) return 0;
  std::lock_guard<std::mutex> lk(s->mu);
  return (uint32_t)s->graphs.size();
}

}  // extern "C"
