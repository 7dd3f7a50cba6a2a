/*
 * Drop-in for DeviceStateProcessingLogic
 * (service-device-state/src/main/java/com/sitewhere/devicestate/processing/DeviceStateProcessingLogic.java:65-201):
 * same SPI (IDeviceStateProcessingLogic.process(List<ConsumerRecord<String, byte[]>>)), same resulting device state.
 *
 * The reference handles the batch record by record: parse the GEnrichedEventPayload, read the device's state
 * document, overwrite lastLocationEventId / lastMeasurementEventIds[name] / lastAlertEventIds[type]
 * (+ lastInteractionDate) and write the document back — one read-modify-write of the store per EVENT.  Here the
 * record values of the whole poll() batch go to the GPU as they are (swgpu_decode_merge with SWGPU_FMT_ENRICHED:
 * the library parses the enriched payload itself and folds it, per (assignment id, kind, name) the newest event
 * wins, exactly what the sequential loop leaves behind), and only the winners are written back: one store update
 * per touched device assignment per flush.
 *
 * Records the library hands back (status 3: command invocations / responses, state changes — events that only
 * matter to the reference when presence was lost, DeviceStateProcessingLogic.java:128-133) go through an embedded
 * stock DeviceStateProcessingLogic, as do none of the others: a record the library rejects (status 1) is one the
 * reference would have logged and skipped (DeviceStateProcessingLogic.java:75-85,101-106).
 *
 * Limit: the GPU table keeps, per (assignment, kind, name), the winner's event id, date and value, and per
 * assignment the presence dates; deviceId / deviceTypeId / customerId / areaId / assetId of the state document
 * (DeviceStateProcessingLogic.java:136-141) are not carried, so a state this shim CREATES holds the assignment id
 * and the merged fields only (an existing document keeps its other fields: null never overwrites,
 * DeviceStatePersistence.java:64-78).
 *
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.nio.charset.StandardCharsets;
import java.util.ArrayList;
import java.util.Date;
import java.util.HashMap;
import java.util.List;
import java.util.Map;
import java.util.UUID;

import org.apache.kafka.clients.consumer.ConsumerRecord;

import com.sitewhere.devicestate.processing.DeviceStateProcessingLogic;
import com.sitewhere.devicestate.spi.microservice.IDeviceStateTenantEngine;
import com.sitewhere.devicestate.spi.processing.IDeviceStateProcessingLogic;
import com.sitewhere.rest.model.device.state.request.DeviceStateCreateRequest;
import com.sitewhere.server.lifecycle.TenantEngineLifecycleComponent;
import com.sitewhere.spi.SiteWhereException;
import com.sitewhere.spi.device.state.IDeviceState;
import com.sitewhere.spi.device.state.IDeviceStateManagement;
import com.sitewhere.spi.server.lifecycle.ILifecycleProgressMonitor;

public class GpuDeviceStateProcessingLogic extends TenantEngineLifecycleComponent implements IDeviceStateProcessingLogic {

    private static final int KIND_MEASUREMENT = 0, KIND_LOCATION = 1, KIND_ALERT = 2, KIND_DEVICE = 0x7f;
    /** SWGPU_SEQ_RECORD_SHIFT of include/swgpu.h */
    private static final int SEQ_RECORD_SHIFT = 12;

    private int deviceOrdinal = 0;
    /** flush the folded state to IDeviceStateManagement every this many batches (1 = after every poll) */
    private int flushEvery = 1;
    private SwGpu.Context ctx;
    private long sequence = 0;
    /** sequence number up to which the store has been brought up to date */
    private long flushedThrough = -1;
    private int sinceFlush = 0;
    private final DeviceStateProcessingLogic stock = new DeviceStateProcessingLogic();

    @Override
    public void start(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	ctx = new SwGpu.Context(deviceOrdinal);
    }

    @Override
    public void stop(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	flush();
	ctx.close();
    }

    @Override
    public void process(List<ConsumerRecord<String, byte[]>> records) throws SiteWhereException {
	int n = records.size();
	if (n == 0) {
	    return;
	}
	long total = 0;
	for (ConsumerRecord<String, byte[]> r : records) {
	    total += r.value().length;
	}
	List<ConsumerRecord<String, byte[]>> handBack = null;
	try (Arena a = Arena.ofConfined()) {
	    MemorySegment bytes = a.allocate(Math.max(total, 1));
	    MemorySegment recOff = a.allocate(JAVA_LONG, n + 1);
	    long off = 0;
	    for (int i = 0; i < n; i++) {
		byte[] v = records.get(i).value();
		recOff.setAtIndex(JAVA_LONG, i, off);
		MemorySegment.copy(v, 0, bytes, JAVA_BYTE, off, v.length);
		off += v.length;
	    }
	    recOff.setAtIndex(JAVA_LONG, n, off);
	    // one native call: parse the enriched payloads on the GPU and fold them into the state; rows never
	    // leave the device.  lastInteractionDate of the touched devices := now (reference: new Date() per event).
	    ctx.check((int) SwGpu.SET_OPTION.invokeExact(ctx.handle, a.allocateUtf8String("state.presence_now_ms"),
		    System.currentTimeMillis()));
	    MemorySegment status = a.allocate(n), outRows = a.allocate(4);
	    ctx.check((int) SwGpu.DECODE_MERGE.invokeExact(ctx.handle, SwGpu.FMT_ENRICHED, bytes, recOff, n, sequence,
		    MemorySegment.NULL, MemorySegment.NULL, status, outRows));
	    sequence += ((long) n) << SEQ_RECORD_SHIFT;
	    for (int i = 0; i < n; i++) {
		if (status.get(JAVA_BYTE, i) == SwGpu.ST_PASSTHROUGH) {
		    if (handBack == null) {
			handBack = new ArrayList<>();
		    }
		    handBack.add(records.get(i));
		}
	    }
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable t) {
	    throw new SiteWhereException(t);
	}
	if (++sinceFlush >= flushEvery || handBack != null) {
	    flush(); // (the stock logic below reads the store: it must see this batch's winners)
	}
	if (handBack != null) {
	    stock.process(handBack);
	}
    }

    /**
     * swgpu_state_snapshot -> one createDeviceState / updateDeviceState per device assignment whose state changed since
     * the last flush, with the fields the reference's merge methods set (DeviceStateProcessingLogic.java:135-159,
     * 169-201; DeviceStatePersistence.deviceStateUpdateLogic, .../persistence/DeviceStatePersistence.java:60-85: ids
     * overwrite, the maps only grow).  A key of the snapshot is token_len(2, LE) | kind | token | name, the token
     * being the 16 bytes of the assignment id.
     */
    void flush() throws SiteWhereException {
	sinceFlush = 0;
	IDeviceStateManagement store = ((IDeviceStateTenantEngine) getTenantEngine()).getDeviceStateManagement();
	Map<UUID, DeviceStateCreateRequest> changed = new HashMap<>();
	try (Arena a = Arena.ofConfined()) {
	    MemorySegment nKeys = a.allocate(JAVA_LONG), keyBytes = a.allocate(JAVA_LONG);
	    ctx.check((int) SwGpu.STATE_SIZE.invokeExact(ctx.handle, nKeys, keyBytes));
	    long n = nKeys.get(JAVA_LONG, 0), kb = keyBytes.get(JAVA_LONG, 0);
	    if (n == 0) {
		return;
	    }
	    MemorySegment seq = a.allocate(JAVA_LONG, n), idHi = a.allocate(JAVA_LONG, n), idLo = a.allocate(JAVA_LONG, n);
	    MemorySegment date = a.allocate(JAVA_LONG, n), value = a.allocate(JAVA_LONG, n), keyOff = a.allocate(JAVA_LONG, n + 1);
	    MemorySegment keys = a.allocate(Math.max(kb, 1));
	    MemorySegment rows = a.allocate(7 * 8); // swgpu_state_rows: seven pointers
	    MemorySegment[] cols = { seq, idHi, idLo, date, value, keyOff, keys };
	    for (int c = 0; c < 7; c++) {
		rows.setAtIndex(java.lang.foreign.ValueLayout.ADDRESS, c, cols[c]);
	    }
	    MemorySegment got = a.allocate(JAVA_LONG);
	    ctx.check((int) SwGpu.STATE_SNAPSHOT.invokeExact(ctx.handle, n, kb, rows, got));
	    n = got.get(JAVA_LONG, 0);
	    long newest = flushedThrough;
	    for (long k = 0; k < n; k++) {
		long o = keyOff.getAtIndex(JAVA_LONG, k), e = keyOff.getAtIndex(JAVA_LONG, k + 1);
		int tl = (keys.get(JAVA_BYTE, o) & 0xff) | ((keys.get(JAVA_BYTE, o + 1) & 0xff) << 8);
		int kind = keys.get(JAVA_BYTE, o + 2) & 0x7f;
		if (tl != 16) {
		    continue; // not an enriched-event key
		}
		UUID assignment = new UUID(keys.get(JAVA_LONG, o + 3), keys.get(JAVA_LONG, o + 11)); // (little-endian words)
		long s = seq.getAtIndex(JAVA_LONG, k);
		DeviceStateCreateRequest req = changed.computeIfAbsent(assignment, id -> {
		    DeviceStateCreateRequest r = new DeviceStateCreateRequest();
		    r.setDeviceAssignmentId(id);
		    return r;
		});
		if (kind == KIND_DEVICE) {
		    // presence row: seq = lastInteractionDate, event_date = presenceMissingDate (0 = null)
		    req.setLastInteractionDate(new Date(s));
		    long missing = date.getAtIndex(JAVA_LONG, k);
		    req.setPresenceMissingDate(missing == 0 ? null : new Date(missing));
		    continue;
		}
		if (s <= flushedThrough) {
		    continue; // unchanged since the last flush
		}
		newest = Math.max(newest, s);
		UUID eventId = new UUID(idHi.getAtIndex(JAVA_LONG, k), idLo.getAtIndex(JAVA_LONG, k));
		byte[] name = new byte[(int) (e - o - 3 - tl)];
		MemorySegment.copy(keys, JAVA_BYTE, o + 3 + tl, name, 0, name.length);
		String nm = new String(name, StandardCharsets.UTF_8);
		if (kind == KIND_LOCATION) {
		    req.setLastLocationEventId(eventId);
		} else if (kind == KIND_MEASUREMENT) {
		    req.getLastMeasurementEventIds().put(nm, eventId);
		} else if (kind == KIND_ALERT) {
		    req.getLastAlertEventIds().put(nm, eventId);
		}
	    }
	    flushedThrough = newest;
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable t) {
	    throw new SiteWhereException(t);
	}
	for (Map.Entry<UUID, DeviceStateCreateRequest> en : changed.entrySet()) {
	    DeviceStateCreateRequest req = en.getValue();
	    if (req.getLastLocationEventId() == null && req.getLastMeasurementEventIds().isEmpty()
		    && req.getLastAlertEventIds().isEmpty()) {
		continue; // only its presence row: nothing newer than the last flush
	    }
	    IDeviceState original = store.getDeviceStateByDeviceAssignmentId(en.getKey());
	    if (original != null) {
		// deviceStateUpdateLogic: null never overwrites, putAll on the maps
		store.updateDeviceState(original.getId(), req);
	    } else {
		store.createDeviceState(req);
	    }
	}
    }

    public void setDeviceOrdinal(int deviceOrdinal) {
	this.deviceOrdinal = deviceOrdinal;
    }

    public void setFlushEvery(int flushEvery) {
	this.flushEvery = flushEvery;
    }
}
