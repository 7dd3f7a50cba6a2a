/*
 * Micro-batcher: IDeviceEventDecoder.decode is one payload per call from many receiver threads
 * (service-event-sources/.../mqtt/MqttInboundEventReceiver.java:175,204-213); the GPU wants a batch.
 * Callers enqueue (payload, future); one flusher thread closes a batch when it reaches maxRecords or
 * lingerMicros, makes ONE native call, and completes every future with that record's outcome.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.util.ArrayList;
import java.util.List;
import java.util.concurrent.ArrayBlockingQueue;
import java.util.concurrent.BlockingQueue;
import java.util.concurrent.CompletableFuture;
import java.util.concurrent.TimeUnit;

import com.sitewhere.spi.SiteWhereException;

final class GpuBatcher implements AutoCloseable {

    /** Columnar view of one decoded row, read by the decoders to rebuild the request POJO. */
    record Row(long meta, long eventDate, double v0, double v1, double v2, long tokenSpan, long nameSpan,
	    long originatorSpan, long messageSpan, long metadataSpan, long alternateIdSpan) {
    }

    /** Outcome for one payload: status + its rows (0..N). */
    record Outcome(int status, List<Row> rows) {
    }

    private record Pending(byte[] payload, CompletableFuture<Outcome> future) {
    }

    private final SwGpu.Context ctx;
    private final int format; // SwGpu.FMT_*
    private final int maxRecords;
    private final long lingerMicros;
    private final BlockingQueue<Pending> queue;
    private final Thread flusher;
    private volatile boolean running = true;

    GpuBatcher(SwGpu.Context ctx, int format, int maxRecords, long lingerMicros) {
	this.ctx = ctx;
	this.format = format;
	this.maxRecords = maxRecords;
	this.lingerMicros = lingerMicros;
	this.queue = new ArrayBlockingQueue<>(maxRecords * 4);
	this.flusher = new Thread(this::run, "swgpu-batcher");
	this.flusher.setDaemon(true);
	this.flusher.start();
    }

    CompletableFuture<Outcome> submit(byte[] payload) throws InterruptedException {
	CompletableFuture<Outcome> f = new CompletableFuture<>();
	queue.put(new Pending(payload, f));
	return f;
    }

    private void run() {
	List<Pending> batch = new ArrayList<>(maxRecords);
	while (running) {
	    try {
		Pending first = queue.poll(50, TimeUnit.MILLISECONDS);
		if (first == null) {
		    continue;
		}
		batch.add(first);
		long deadline = System.nanoTime() + lingerMicros * 1000L;
		while (batch.size() < maxRecords) {
		    long left = deadline - System.nanoTime();
		    if (left <= 0) {
			break;
		    }
		    Pending p = queue.poll(left, TimeUnit.NANOSECONDS);
		    if (p == null) {
			break;
		    }
		    batch.add(p);
		}
		flush(batch);
	    } catch (Throwable t) {
		for (Pending p : batch) {
		    p.future().completeExceptionally(t);
		}
	    } finally {
		batch.clear();
	    }
	}
    }

    /** One native decode call for the whole batch (swgpu_decode_pb / swgpu_decode_json). */
    private void flush(List<Pending> batch) throws Throwable {
	int n = batch.size();
	long total = 0;
	for (Pending p : batch) {
	    total += p.payload().length;
	}
	try (Arena a = Arena.ofConfined()) {
	    // (a production build keeps these segments in swgpu_alloc_pinned memory and reuses them)
	    MemorySegment bytes = a.allocate(Math.max(total, 1));
	    MemorySegment off = a.allocate(JAVA_LONG, n + 1L);
	    long pos = 0;
	    for (int i = 0; i < n; i++) {
		off.setAtIndex(JAVA_LONG, i, pos);
		MemorySegment.copy(batch.get(i).payload(), 0, bytes, JAVA_BYTE, pos, batch.get(i).payload().length);
		pos += batch.get(i).payload().length;
	    }
	    off.setAtIndex(JAVA_LONG, n, pos);
	    int maxRows = n;
	    for (;;) {
		MemorySegment[] col = new MemorySegment[13];
		MemorySegment cols = a.allocate(SwGpu.COLS_BYTES);
		for (int c = 0; c < 13; c++) {
		    col[c] = a.allocate(JAVA_LONG, Math.max(maxRows, 1));
		    cols.setAtIndex(ADDRESS, c, col[c]);
		}
		MemorySegment status = a.allocate(Math.max(n, 1));
		MemorySegment firstRow = a.allocate(JAVA_INT, n + 1L);
		MemorySegment outRows = a.allocate(JAVA_INT);
		var call = format == SwGpu.FMT_JSON ? SwGpu.DECODE_JSON : SwGpu.DECODE_PB;
		int rc = (int) call.invokeExact(ctx.handle, bytes, off, n, maxRows, cols, status, firstRow, outRows);
		if (rc == SwGpu.E_CAPACITY) { // measurement fan-out: retry with the exact row count
		    maxRows = outRows.get(JAVA_INT, 0);
		    continue;
		}
		ctx.check(rc);
		for (int i = 0; i < n; i++) {
		    int r0 = firstRow.getAtIndex(JAVA_INT, i), r1 = firstRow.getAtIndex(JAVA_INT, i + 1);
		    List<Row> rows = new ArrayList<>(r1 - r0);
		    for (int r = r0; r < r1; r++) {
			rows.add(new Row(col[0].getAtIndex(JAVA_LONG, r), col[3].getAtIndex(JAVA_LONG, r),
				col[4].getAtIndex(JAVA_DOUBLE, r), col[5].getAtIndex(JAVA_DOUBLE, r),
				col[6].getAtIndex(JAVA_DOUBLE, r), col[7].getAtIndex(JAVA_LONG, r),
				col[8].getAtIndex(JAVA_LONG, r), col[9].getAtIndex(JAVA_LONG, r),
				col[10].getAtIndex(JAVA_LONG, r), col[11].getAtIndex(JAVA_LONG, r),
				col[12].getAtIndex(JAVA_LONG, r)));
		    }
		    batch.get(i).future().complete(new Outcome(status.get(JAVA_BYTE, i), rows));
		}
		return;
	    }
	}
    }

    @Override
    public void close() throws SiteWhereException {
	running = false;
	flusher.interrupt();
    }
}
