package co.fanki.domainmcp.analysis.domain;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemoryLayout;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.StructLayout;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;

/**
 * Downcall handles for libdmgraph.so (include/dmgraph.h). NOT COMPILED in the dmgraph repository
 * (no JDK in its build image); shipped as the binding a maintainer adds to the host.
 */
final class DmGraphNative {

    static final int OK = 0, EINVAL = 1, ENOMEM = 2, ECAP = 3, ECUDA = 4, ENCCL = 5, ENODEV = 6;
    static final int OUT = 0, IN = 1, BOTH = 2;
    static final int SEARCH_CONTAINS_ALL = 0, SEARCH_SIMPLE_EQ_FIRST = 1, SEARCH_CONTAINS_FIRST = 2;

    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB = SymbolLookup.libraryLookup("libdmgraph.so", Arena.global());

    /** dmg_graph_desc: 4 x uint32, uint64, 7 pointers (include/dmgraph.h). */
    static final StructLayout GRAPH_DESC = MemoryLayout.structLayout(
            JAVA_INT.withName("n_vertices"), JAVA_INT.withName("n_entry"),
            JAVA_INT.withName("n_parts"), JAVA_INT.withName("reserved"),
            JAVA_LONG.withName("n_edges"),
            ADDRESS.withName("name_off"), ADDRESS.withName("name_bytes"), ADDRESS.withName("lower_bytes"),
            ADDRESS.withName("edge_src"), ADDRESS.withName("edge_dst"),
            ADDRESS.withName("entry_ids"), ADDRESS.withName("part_off"));

    private static MethodHandle h(final String name, final FunctionDescriptor fd) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(), fd);
    }

    static final MethodHandle ENGINE_CREATE = h("dmg_engine_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle ENGINE_DESTROY = h("dmg_engine_destroy", FunctionDescriptor.ofVoid(ADDRESS));
    static final MethodHandle LAST_ERROR = h("dmg_last_error", FunctionDescriptor.of(ADDRESS));
    static final MethodHandle GRAPH_CREATE = h("dmg_graph_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle GRAPH_RETAIN = h("dmg_graph_retain", FunctionDescriptor.ofVoid(ADDRESS));
    static final MethodHandle GRAPH_RELEASE = h("dmg_graph_release", FunctionDescriptor.ofVoid(ADDRESS));
    static final MethodHandle ADJACENCY_BATCH = h("dmg_adjacency_batch", FunctionDescriptor.of(
            JAVA_INT, ADDRESS, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    static final MethodHandle NEIGHBORS_BATCH = h("dmg_neighbors_batch", FunctionDescriptor.of(
            JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS,
            ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    static final MethodHandle SEARCH_BATCH = h("dmg_search_batch", FunctionDescriptor.of(
            JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS,
            ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    static final MethodHandle ANALYSIS_ORDER = h("dmg_analysis_order", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle LOOKUP_BATCH = h("dmg_lookup_batch", FunctionDescriptor.of(
            JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS));

    /** Status to exception, the way the reference reports errors. */
    static void check(final int status) {
        if (status == OK) {
            return;
        }
        String msg;
        try {
            msg = ((MemorySegment) LAST_ERROR.invokeExact()).reinterpret(512).getUtf8String(0);
        } catch (final Throwable t) {
            msg = "dmgraph status " + status;
        }
        if (status == EINVAL) {
            throw new IllegalArgumentException(msg);      // shared/Preconditions.java:41-47
        }
        throw new IllegalStateException(msg);
    }

    private DmGraphNative() {
    }
}
