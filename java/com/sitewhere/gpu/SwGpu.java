/*
 * Panama (java.lang.foreign) binding of libswgpu.so — see include/swgpu.h for the contract of each call.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;

import com.sitewhere.spi.SiteWhereException;

public final class SwGpu {

    public static final int ST_OK = 0, ST_DECODE_EXCEPTION = 1, ST_RUNTIME_EXCEPTION = 2, ST_PASSTHROUGH = 3;
    public static final int FMT_PROTOBUF = 0, FMT_JSON = 1, FMT_ENRICHED = 2;
    public static final int E_CAPACITY = -4;
    /** swgpu_decoded_cols: 13 pointers (9 mandatory, 4 optional). */
    public static final long COLS_BYTES = 13 * 8;

    private static final Linker L = Linker.nativeLinker();
    private static final SymbolLookup LIB = SymbolLookup.libraryLookup(System.getProperty("swgpu.lib", "libswgpu.so"),
	    Arena.global());

    private static MethodHandle h(String name, FunctionDescriptor d) {
	return L.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
    }

    static final MethodHandle CREATE = h("swgpu_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle DESTROY = h("swgpu_destroy", FunctionDescriptor.ofVoid(ADDRESS));
    static final MethodHandle LAST_ERROR = h("swgpu_last_error", FunctionDescriptor.of(ADDRESS, ADDRESS));
    static final MethodHandle ALLOC_PINNED = h("swgpu_alloc_pinned", FunctionDescriptor.of(ADDRESS, JAVA_LONG));
    static final MethodHandle FREE_PINNED = h("swgpu_free_pinned", FunctionDescriptor.ofVoid(ADDRESS));
    static final MethodHandle SET_ZONES = h("swgpu_set_zones",
	    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle DECODE_PB = h("swgpu_decode_pb", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
	    JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle DECODE_JSON = h("swgpu_decode_json", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS,
	    ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle ZONE_TEST = h("swgpu_zone_test", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
	    ADDRESS, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS));
    static final MethodHandle DECODE_ZONE_TEST = h("swgpu_decode_zone_test",
	    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS,
		    ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, JAVA_INT));
    static final MethodHandle MERGE = h("swgpu_merge", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
	    JAVA_INT, ADDRESS, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS));
    static final MethodHandle DECODE_MERGE = h("swgpu_decode_merge", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
	    ADDRESS, ADDRESS, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle STATE_SIZE = h("swgpu_state_size", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle STATE_SNAPSHOT = h("swgpu_state_snapshot",
	    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS));
    static final MethodHandle STATE_RESTORE = h("swgpu_state_restore", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG));
    static final MethodHandle SET_OPTION = h("swgpu_set_option", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG));
    static final MethodHandle RESERVE_BATCH = h("swgpu_reserve_batch", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_LONG));
    static final MethodHandle DECODE_ENRICHED = h("swgpu_decode_enriched", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS,
	    ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle SUBMIT_DECODE_ZONE_TEST = h("swgpu_submit_decode_zone_test",
	    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS,
		    ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, JAVA_INT, ADDRESS));
    static final MethodHandle WAIT = h("swgpu_wait", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG));
    static final MethodHandle PRESENCE_SCAN = h("swgpu_presence_scan", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG,
	    JAVA_LONG, JAVA_LONG, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    /** device-buffer forms, for callers that stage their own buffers (pointers are CUdeviceptr values) */
    static final MethodHandle DECODE_ZONE_TEST_DEV = h("swgpu_decode_zone_test_dev",
	    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS,
		    ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
    /** the GInboundEventPayload bytes of every decoded row (EventSourcesManager.handleDecodedEvent's marshal step) */
    static final MethodHandle ENCODE_INBOUND_DEV = h("swgpu_encode_inbound_dev",
	    FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_INT,
		    JAVA_LONG, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS));
    /** AlternateIdDeduplicator.isDuplicate as a set probe: dup[row] = 0 new, 1 seen, 2 ask the store */
    static final MethodHandle DEDUP_PROBE_DEV = h("swgpu_dedup_probe_dev", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS,
	    ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle DEDUP_PROBE = h("swgpu_dedup_probe", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
	    JAVA_INT, ADDRESS));
    static final MethodHandle PARTITION = h("swgpu_partition", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT));

    /** One native context: one GPU, one calling thread. */
    public static final class Context implements AutoCloseable {
	final MemorySegment handle;

	public Context(int deviceOrdinal) throws SiteWhereException {
	    try (Arena a = Arena.ofConfined()) {
		MemorySegment out = a.allocate(ADDRESS);
		int rc = (int) CREATE.invokeExact(deviceOrdinal, MemorySegment.NULL, out);
		if (rc != 0) {
		    throw new SiteWhereException("swgpu_create failed: " + lastError(MemorySegment.NULL));
		}
		this.handle = out.get(ADDRESS, 0);
	    } catch (SiteWhereException e) {
		throw e;
	    } catch (Throwable t) {
		throw new SiteWhereException(t);
	    }
	}

	void check(int rc) throws SiteWhereException {
	    if (rc != 0) {
		throw new SiteWhereException("libswgpu error " + rc + ": " + lastError(handle));
	    }
	}

	@Override
	public void close() {
	    try {
		DESTROY.invokeExact(handle);
	    } catch (Throwable ignored) {
	    }
	}
    }

    static String lastError(MemorySegment ctx) {
	try {
	    MemorySegment p = (MemorySegment) LAST_ERROR.invokeExact(ctx);
	    return p.reinterpret(512).getString(0);
	} catch (Throwable t) {
	    return "?";
	}
    }

    /** Pinned staging the batcher fills with payload bytes (wrapped as a MemorySegment). */
    static MemorySegment pinned(long bytes) throws SiteWhereException {
	try {
	    MemorySegment p = (MemorySegment) ALLOC_PINNED.invokeExact(bytes);
	    if (p.equals(MemorySegment.NULL)) {
		throw new SiteWhereException("swgpu_alloc_pinned(" + bytes + ") failed");
	    }
	    return p.reinterpret(bytes);
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable t) {
	    throw new SiteWhereException(t);
	}
    }

    private SwGpu() {
    }
}
