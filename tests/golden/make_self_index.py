#!/usr/bin/env python3
"""BASELINE config 1 fixture: the reference indexing ITSELF.

Walks /root/reference/src/main/java the way SourceParser.parse (SourceParser.java:146-194)
and JavaSourceParser (JavaSourceParser.java:119-189, 199-223, 539-564) do and writes the
resulting graph -- identifiers, import edges, entry points -- to self_index.json.

  * files: every *.java under src/main/java, sorted by path       (JavaSourceParser.java:129-135)
  * id:    path relative to the source root, '/'->'.', no '.java' (JavaSourceParser.java:149-160)
  * edges: explicit imports; wildcard imports skipped; static imports cut back to the
           class; kept only if the target is a known id           (JavaSourceParser.java:174-189, 539-564)
  * the per-file dependency set is a java.util.HashSet<String>, so the order in which
    addDependency sees the targets is HashSet iteration order; reproduced here with an
    emulation of String.hashCode + HashMap bucket order (JDK 21 semantics)
  * entry points: a type- or method-level annotation named RestController, Controller,
    KafkaListener, Scheduled, EventListener or SpringBootApplication
                                                                  (JavaSourceParser.java:72-79, 199-223)

Only names and edges are written -- no reference source text.  This script needs
/root/reference and therefore runs only in the build container; the JSON it writes is
committed and is what the tests (and the GPU box) read.
"""
import json
import os
import re
import sys

REF = "/root/reference"
SRC = os.path.join(REF, "src/main/java")
ENTRY = {"RestController", "Controller", "KafkaListener", "Scheduled", "EventListener",
         "SpringBootApplication"}


def java_string_hash(s):
    h = 0
    for ch in s.encode("utf-16-be"):
        pass
    units = s.encode("utf-16-be")
    for i in range(0, len(units), 2):
        c = (units[i] << 8) | units[i + 1]
        h = (31 * h + c) & 0xFFFFFFFF
    return h


def java_hashset_order(items):
    """Iteration order of a java.util.HashSet<String> after add()-ing `items` in order.

    HashMap: spread = h ^ (h >>> 16); bucket = spread & (cap-1); cap starts at 16 and
    doubles when size > 0.75*cap; a resize splits each bucket preserving relative order;
    iteration walks buckets ascending, each chain in insertion order.  Chains never reach
    the treeify threshold (8) for the set sizes that occur here; asserted below.
    """
    cap = 16
    buckets = [[] for _ in range(cap)]
    seen = set()
    size = 0

    def spread(s):
        h = java_string_hash(s)
        return (h ^ (h >> 16)) & 0xFFFFFFFF

    for it in items:
        if it in seen:
            continue
        seen.add(it)
        buckets[spread(it) & (cap - 1)].append(it)
        size += 1
        if size > 0.75 * cap:
            ncap = cap * 2
            nb = [[] for _ in range(ncap)]
            for b in buckets:            # ascending old bucket; relative order preserved
                for x in b:
                    nb[spread(x) & (ncap - 1)].append(x)
            cap, buckets = ncap, nb
    assert all(len(b) < 8 for b in buckets), "treeified bin: order emulation not valid"
    return [x for b in buckets for x in b]


def strip_comments(text):
    out = []
    i, n = 0, len(text)
    while i < n:
        c = text[i]
        if text.startswith('"""', i):
            j = text.find('"""', i + 3)
            j = n if j < 0 else j + 3
            out.append('""')
            i = j
        elif c == '"':
            j = i + 1
            while j < n and text[j] != '"':
                j += 2 if text[j] == "\\" else 1
            out.append('""')
            i = j + 1
        elif c == "'":
            j = i + 1
            while j < n and text[j] != "'":
                j += 2 if text[j] == "\\" else 1
            out.append("''")
            i = j + 1
        elif text.startswith("//", i):
            j = text.find("\n", i)
            i = n if j < 0 else j
        elif text.startswith("/*", i):
            j = text.find("*/", i + 2)
            i = n if j < 0 else j + 2
            out.append(" ")
        else:
            out.append(c)
            i += 1
    return "".join(out)


IMPORT_RE = re.compile(r"^\s*import\s+(static\s+)?([A-Za-z_][\w.]*?)(\.\*)?\s*;", re.M)
ANN_RE = re.compile(r"@([A-Za-z_][\w.]*)")


def main():
    if not os.path.isdir(SRC):
        sys.exit("needs %s (build container only)" % SRC)
    files = []
    for root, _dirs, names in os.walk(SRC):
        for nm in names:
            if nm.endswith(".java"):
                files.append(os.path.join(root, nm))
    files.sort()   # java.nio.file.Path natural order on unix == byte order of the path string

    ids = []
    for f in files:
        rel = os.path.relpath(f, SRC)
        ids.append(rel[:-5].replace("/", "."))
    known = set(ids)

    nodes, edges, entries = [], [], []
    for f, ident in zip(files, ids):
        nodes.append([ident, os.path.relpath(f, REF)])
    for f, ident in zip(files, ids):
        text = strip_comments(open(f, encoding="utf-8").read())
        # imports precede the first type declaration
        m = re.search(r"\b(class|interface|enum|record)\s+[A-Z]", text)
        head = text[: m.start()] if m else text
        imported = []
        for im in IMPORT_RE.finditer(head):
            is_static, name, star = im.group(1), im.group(2), im.group(3)
            if star:
                continue
            if is_static:
                k = name.rfind(".")
                if k <= 0:
                    continue
                name = name[:k]
            if name in known:
                imported.append(name)
        for tgt in java_hashset_order(imported):
            edges.append([ident, tgt])
    for f, ident in zip(files, ids):
        text = strip_comments(open(f, encoding="utf-8").read())
        anns = {a.split(".")[-1] for a in ANN_RE.findall(text)}
        if anns & ENTRY:
            entries.append(ident)

    out = {
        "about": "waabox/domain-mcp-server src/main/java indexed by JavaSourceParser's rules; "
                 "made by tests/golden/make_self_index.py",
        "nodes": nodes, "edges": edges, "entryPoints": entries,
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "self_index.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1)
    print("wrote %s: %d nodes, %d edges, %d entry points" % (path, len(nodes), len(edges), len(entries)))


if __name__ == "__main__":
    main()
