/*
 * Drop-in for DeviceStateProcessingLogic
 * (service-device-state/src/main/java/com/sitewhere/devicestate/processing/DeviceStateProcessingLogic.java:65-201):
 * same SPI (IDeviceStateProcessingLogic.process(List<ConsumerRecord<String, byte[]>>)), same resulting device state.
 *
 * The reference handles the batch record by record: unmarshal the enriched event, read the device's state document,
 * overwrite lastLocationEventId / lastMeasurementEventIds[name] / lastAlertEventIds[type] (+ lastInteractionDate) and
 * write the document back — one read-modify-write of the store per EVENT.  Here the whole poll() batch is folded on
 * the GPU first (swgpu_decode_merge: per (device, kind, name) the newest event wins, exactly what the sequential loop leaves
 * behind), and only the winners are written: one store update per touched DEVICE per flush.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.util.List;

import org.apache.kafka.clients.consumer.ConsumerRecord;

import com.sitewhere.devicestate.spi.processing.IDeviceStateProcessingLogic;
import com.sitewhere.server.lifecycle.TenantEngineLifecycleComponent;
import com.sitewhere.spi.SiteWhereException;
import com.sitewhere.spi.server.lifecycle.ILifecycleProgressMonitor;

public class GpuDeviceStateProcessingLogic extends TenantEngineLifecycleComponent implements IDeviceStateProcessingLogic {

    private int deviceOrdinal = 0;
    /** flush the folded state to IDeviceStateManagement every this many batches (1 = after every poll) */
    private int flushEvery = 1;
    private SwGpu.Context ctx;
    private long sequence = 0;
    private int sinceFlush = 0;

    @Override
    public void start(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	ctx = new SwGpu.Context(deviceOrdinal);
    }

    @Override
    public void stop(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	flush();
	ctx.close();
    }

    /*
     * The enriched-event topic carries protobuf-marshaled payloads (EventModelMarshaler); the event part re-encodes
     * to the same DeviceEvent wire format the inbound decoder reads, so the batch is staged as
     * (bytes, rec_off) + the 128-bit event ids, decoded and merged without leaving the GPU.
     */
    @Override
    public void process(List<ConsumerRecord<String, byte[]>> records) throws SiteWhereException {
	int n = records.size();
	if (n == 0) {
	    return;
	}
	long total = 0;
	for (ConsumerRecord<String, byte[]> r : records) {
	    total += EnrichedEvents.eventBytes(r.value()).length;
	}
	try (Arena a = Arena.ofConfined()) {
	    MemorySegment bytes = a.allocate(Math.max(total, 1));
	    MemorySegment recOff = a.allocate(JAVA_LONG, n + 1);
	    MemorySegment idHi = a.allocate(JAVA_LONG, n), idLo = a.allocate(JAVA_LONG, n);
	    long off = 0;
	    for (int i = 0; i < n; i++) {
		byte[] ev = EnrichedEvents.eventBytes(records.get(i).value());
		recOff.setAtIndex(JAVA_LONG, i, off);
		MemorySegment.copy(ev, 0, bytes, JAVA_BYTE, off, ev.length);
		off += ev.length;
		java.util.UUID id = EnrichedEvents.eventId(records.get(i).value());
		idHi.setAtIndex(JAVA_LONG, i, id.getMostSignificantBits());
		idLo.setAtIndex(JAVA_LONG, i, id.getLeastSignificantBits());
	    }
	    recOff.setAtIndex(JAVA_LONG, n, off);
	    // one native call: decode on the GPU and fold into the state, rows never leave the device
	    MemorySegment status = a.allocate(n), outRows = a.allocate(4);
	    ctx.check((int) SwGpu.DECODE_MERGE.invokeExact(ctx.handle, SwGpu.FMT_PROTOBUF, bytes, recOff, n, sequence, idHi, idLo,
		    status, outRows));
	    sequence += outRows.get(java.lang.foreign.ValueLayout.JAVA_INT, 0);
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable t) {
	    throw new SiteWhereException(t);
	}
	if (++sinceFlush >= flushEvery) {
	    flush();
	}
    }

    /**
     * swgpu_state_snapshot -> one createDeviceState / updateDeviceState per device, with the same fields the
     * reference's merge methods set (DeviceStateProcessingLogic.java:165-201).
     */
    void flush() throws SiteWhereException {
	sinceFlush = 0;
	StateWriter.write(ctx, getTenantEngine());
    }

    public void setDeviceOrdinal(int deviceOrdinal) {
	this.deviceOrdinal = deviceOrdinal;
    }

    public void setFlushEvery(int flushEvery) {
	this.flushEvery = flushEvery;
    }
}
