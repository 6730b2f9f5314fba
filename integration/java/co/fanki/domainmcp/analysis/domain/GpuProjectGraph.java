package co.fanki.domainmcp.analysis.domain;

import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.nio.charset.StandardCharsets;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.LinkedHashSet;
import java.util.List;
import java.util.Map;
import java.util.Set;

/**
 * GPU-side twin of one published {@link ProjectGraph}. NOT COMPILED in the dmgraph repository.
 *
 * <p>{@code ProjectGraph} is {@code public final} (ProjectGraph.java:34), so this class does not extend
 * it: it is a delegate the graph holds. The host edit is confined to ProjectGraph itself -- one field,
 * one setter and a first line in each of the four read methods (see README.md in this directory):</p>
 * <pre>
 *   private GpuProjectGraph gpu;                                   // null until GraphService publishes
 *   void attachGpu(GpuProjectGraph g) { this.gpu = g; }
 *   public Set&lt;String&gt; dependencies(String id) { if (gpu != null) return gpu.dependencies(id); ...as before... }
 * </pre>
 * <p>Build methods (addNode, addDependency, markAsEntryPoint, metadata) stay the Java maps;
 * {@link #publish} freezes the graph, uploads it (ProjectGraph.java:160-197 rules are re-applied by
 * dmg_graph_create) and from then on dependencies / dependents / resolve / analysisOrder and the
 * identifier scans are single calls into libdmgraph.so. A published graph is immutable, exactly like
 * every graph GraphService holds (GraphService.java:118-122).</p>
 */
public final class GpuProjectGraph implements AutoCloseable {

    private final MemorySegment handle;          // dmg_graph*
    private final String[] names;                // vertex id -> identifier (nodes order)
    private final Map<String, Integer> index;    // identifier -> vertex id

    private GpuProjectGraph(final MemorySegment theHandle, final String[] theNames,
            final Map<String, Integer> theIndex) {
        this.handle = theHandle;
        this.names = theNames;
        this.index = theIndex;
    }

    /** Uploads {@code graph}; called from GraphService.put / reload / loadAll. */
    public static GpuProjectGraph publish(final MemorySegment engine, final ProjectGraph graph) {
        final String[] names = graph.identifiers().toArray(new String[0]);
        final Map<String, Integer> index = new HashMap<>(names.length * 2);
        for (int i = 0; i < names.length; i++) {
            index.put(names[i], i);
        }
        try (Arena arena = Arena.ofConfined()) {
            // name arena + the host's own toLowerCase() copy (CodeContextService.java:1513)
            final MemorySegment nameOff = arena.allocate(JAVA_LONG, names.length + 1L);
            final byte[][] raw = new byte[names.length][];
            final byte[][] low = new byte[names.length][];
            long total = 0;
            for (int i = 0; i < names.length; i++) {
                raw[i] = names[i].getBytes(StandardCharsets.UTF_8);
                low[i] = names[i].toLowerCase().getBytes(StandardCharsets.UTF_8);
                if (low[i].length != raw[i].length) {
                    low[i] = raw[i];             // length-changing folds: fall back to the raw bytes
                }
                nameOff.setAtIndex(JAVA_LONG, i, total);
                total += raw[i].length;
            }
            nameOff.setAtIndex(JAVA_LONG, names.length, total);
            final MemorySegment nameBytes = arena.allocate(Math.max(total, 1));
            final MemorySegment lowerBytes = arena.allocate(Math.max(total, 1));
            long at = 0;
            for (int i = 0; i < names.length; i++) {
                MemorySegment.copy(raw[i], 0, nameBytes, JAVA_BYTE, at, raw[i].length);
                MemorySegment.copy(low[i], 0, lowerBytes, JAVA_BYTE, at, low[i].length);
                at += raw[i].length;
            }
            // edges in dependencies() order per source, entry points in insertion order
            final List<int[]> edges = new ArrayList<>();
            for (final String from : names) {
                for (final String to : graph.dependencies(from)) {
                    edges.add(new int[] {index.get(from), index.getOrDefault(to, -1)});
                }
            }
            final MemorySegment src = arena.allocate(JAVA_INT, Math.max(edges.size(), 1));
            final MemorySegment dst = arena.allocate(JAVA_INT, Math.max(edges.size(), 1));
            for (int i = 0; i < edges.size(); i++) {
                src.setAtIndex(JAVA_INT, i, edges.get(i)[0]);
                dst.setAtIndex(JAVA_INT, i, edges.get(i)[1]);
            }
            final Set<String> eps = graph.entryPoints();
            final MemorySegment entry = arena.allocate(JAVA_INT, Math.max(eps.size(), 1));
            int k = 0;
            for (final String ep : eps) {
                entry.setAtIndex(JAVA_INT, k++, index.get(ep));
            }
            final MemorySegment desc = arena.allocate(DmGraphNative.GRAPH_DESC);
            desc.set(JAVA_INT, 0, names.length);
            desc.set(JAVA_INT, 4, eps.size());
            desc.set(JAVA_INT, 8, 0);
            desc.set(JAVA_LONG, 16, edges.size());
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 24, nameOff);
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 32, nameBytes);
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 40, lowerBytes);
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 48, src);
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 56, dst);
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 64, entry);
            desc.set(java.lang.foreign.ValueLayout.ADDRESS, 72, MemorySegment.NULL);
            final MemorySegment out = arena.allocate(java.lang.foreign.ValueLayout.ADDRESS);
            DmGraphNative.check((int) DmGraphNative.GRAPH_CREATE.invokeExact(engine, desc, out));
            return new GpuProjectGraph(out.get(java.lang.foreign.ValueLayout.ADDRESS, 0), names, index);
        } catch (final RuntimeException e) {
            throw e;
        } catch (final Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /** ProjectGraph.java:299-305. */
    public Set<String> dependencies(final String identifier) {
        return row(DmGraphNative.OUT, identifier);
    }

    /** ProjectGraph.java:316-328 (no scan of every source any more). */
    public Set<String> dependents(final String identifier) {
        return row(DmGraphNative.IN, identifier);
    }

    /** ProjectGraph.java:359-378: dmg_neighbors_batch(BOTH, k = 1, exclude_sources = 0). */
    public Set<String> resolve(final String identifier) {
        final Integer v = identifier == null ? null : index.get(identifier);
        if (v == null) {
            return Set.of();
        }
        try (Arena arena = Arena.ofConfined()) {
            final MemorySegment qOff = arena.allocateArray(JAVA_INT, 0, 1);
            final MemorySegment qSrc = arena.allocateArray(JAVA_INT, v);
            final MemorySegment outOff = arena.allocate(JAVA_LONG, 2);
            final MemorySegment needed = arena.allocate(JAVA_LONG);
            int cap = 256;
            while (true) {
                final MemorySegment ids = arena.allocate(JAVA_INT, cap);
                final MemorySegment hop = arena.allocate(cap);
                final int st = (int) DmGraphNative.NEIGHBORS_BATCH.invokeExact(handle, DmGraphNative.BOTH, 1, 0,
                        1L, qOff, qSrc, outOff, ids, hop, (long) cap, needed);
                if (st == DmGraphNative.ECAP) {
                    cap = (int) needed.get(JAVA_LONG, 0);
                    continue;
                }
                DmGraphNative.check(st);
                return toSet(ids, (int) needed.get(JAVA_LONG, 0));
            }
        } catch (final RuntimeException e) {
            throw e;
        } catch (final Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /** ProjectGraph.java:390-421, exact order. */
    public List<String> analysisOrder() {
        if (names.length == 0) {
            return List.of();
        }
        try (Arena arena = Arena.ofConfined()) {
            final MemorySegment order = arena.allocate(JAVA_INT, names.length);
            DmGraphNative.check((int) DmGraphNative.ANALYSIS_ORDER.invokeExact(handle, order));
            final List<String> out = new ArrayList<>(names.length);
            for (int i = 0; i < names.length; i++) {
                out.add(names[order.getAtIndex(JAVA_INT, i)]);
            }
            return List.copyOf(out);
        } catch (final RuntimeException e) {
            throw e;
        } catch (final Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /**
     * The identifier scan of CodeContextService.searchProject (CodeContextService.java:1507-1526):
     * ids of every identifier whose lower-cased form contains the lower-cased query, nodes order.
     */
    public List<String> searchIdentifiers(final String query) {
        final byte[] q = query.toLowerCase().getBytes(StandardCharsets.UTF_8);   // NPE on null, as in Java
        try (Arena arena = Arena.ofConfined()) {
            final MemorySegment qOff = arena.allocateArray(JAVA_INT, 0, q.length);
            final MemorySegment qBytes = arena.allocateArray(JAVA_BYTE, q.length == 0 ? new byte[1] : q);
            final MemorySegment outOff = arena.allocate(JAVA_LONG, 2);
            final MemorySegment needed = arena.allocate(JAVA_LONG);
            final MemorySegment ids = arena.allocate(JAVA_INT, Math.max(names.length, 1));
            DmGraphNative.check((int) DmGraphNative.SEARCH_BATCH.invokeExact(handle,
                    DmGraphNative.SEARCH_CONTAINS_ALL, 0, names.length, 1L, qOff, qBytes, outOff, ids,
                    (long) names.length, needed));
            final int n = (int) needed.get(JAVA_LONG, 0);
            final List<String> out = new ArrayList<>(n);
            for (int i = 0; i < n; i++) {
                out.add(names[ids.getAtIndex(JAVA_INT, i)]);
            }
            return out;
        } catch (final RuntimeException e) {
            throw e;
        } catch (final Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    private Set<String> row(final int dir, final String identifier) {
        final Integer v = identifier == null ? null : index.get(identifier);
        if (v == null) {
            return Set.of();                      // ProjectGraph.java:300-302, 317-319
        }
        try (Arena arena = Arena.ofConfined()) {
            final MemorySegment id = arena.allocateArray(JAVA_INT, v);
            final MemorySegment outOff = arena.allocate(JAVA_LONG, 2);
            final MemorySegment needed = arena.allocate(JAVA_LONG);
            int cap = 64;
            while (true) {
                final MemorySegment ids = arena.allocate(JAVA_INT, cap);
                final int st = (int) DmGraphNative.ADJACENCY_BATCH.invokeExact(handle, dir, 1L, id, outOff, ids,
                        (long) cap, needed);
                if (st == DmGraphNative.ECAP) {
                    cap = (int) needed.get(JAVA_LONG, 0);
                    continue;
                }
                DmGraphNative.check(st);
                return toSet(ids, (int) needed.get(JAVA_LONG, 0));
            }
        } catch (final RuntimeException e) {
            throw e;
        } catch (final Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    private Set<String> toSet(final MemorySegment ids, final int n) {
        final Set<String> out = new LinkedHashSet<>(n * 2);
        for (int i = 0; i < n; i++) {
            out.add(names[ids.getAtIndex(JAVA_INT, i)]);
        }
        return java.util.Collections.unmodifiableSet(out);
    }

    @Override
    public void close() {
        try {
            DmGraphNative.GRAPH_RELEASE.invokeExact(handle);
        } catch (final Throwable t) {
            throw new IllegalStateException(t);
        }
    }
}
