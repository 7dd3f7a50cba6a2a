/*
 * Drop-in for AlternateIdDeduplicator
 * (service-event-sources/src/main/java/com/sitewhere/sources/deduplicator/AlternateIdDeduplicator.java:42-57):
 * same SPI (IDeviceEventDeduplicator.isDuplicate(IDecodedDeviceRequest), called once per decoded request at
 * InboundEventSource.java:226-227), same answer.
 *
 * The reference asks the event store for every request that carries an alternateId
 * (getDeviceEventByAlternateId: a gRPC round trip + an index lookup).  Here the GPU keeps the set of alternate ids
 * this event source has been shown (swgpu_dedup_probe).  "Seen" (verdict 1) is answered at once: such an id belongs to
 * a request this source already forwarded, i.e. to an event that is stored or on its way to the store.  "New" (0) and
 * "ask" (2: ids over 72 bytes) still go to the stock deduplicator, so an id stored by ANOTHER source or before this
 * process started is found exactly as today; its answer is what is returned.
 *
 * Limit, stated: an event this source forwarded but the pipeline later dropped (failed persistence) is reported as a
 * duplicate when its id comes again, where the reference would let the retry through.  Deployments that rely on that
 * keep the stock bean.
 *
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.nio.charset.StandardCharsets;

import com.sitewhere.server.lifecycle.TenantEngineLifecycleComponent;
import com.sitewhere.sources.deduplicator.AlternateIdDeduplicator;
import com.sitewhere.sources.spi.IDecodedDeviceRequest;
import com.sitewhere.sources.spi.IDeviceEventDeduplicator;
import com.sitewhere.spi.SiteWhereException;
import com.sitewhere.spi.device.event.request.IDeviceEventCreateRequest;
import com.sitewhere.spi.server.lifecycle.ILifecycleProgressMonitor;
import com.sitewhere.spi.server.lifecycle.LifecycleComponentType;

public class GpuAlternateIdDeduplicator extends TenantEngineLifecycleComponent implements IDeviceEventDeduplicator {

    private int deviceOrdinal = 0;
    private SwGpu.Context ctx;
    private final AlternateIdDeduplicator stock = new AlternateIdDeduplicator();

    public GpuAlternateIdDeduplicator() {
	super(LifecycleComponentType.DeviceEventDeduplicator);
    }

    @Override
    public void start(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	ctx = new SwGpu.Context(deviceOrdinal);
    }

    @Override
    public void stop(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	ctx.close();
    }

    @Override
    public synchronized boolean isDuplicate(IDecodedDeviceRequest<?> request) throws SiteWhereException {
	if (!(request.getRequest() instanceof IDeviceEventCreateRequest)) {
	    return false;
	}
	String alternateId = ((IDeviceEventCreateRequest) request.getRequest()).getAlternateId();
	if (alternateId == null) {
	    return false;
	}
	byte[] id = alternateId.getBytes(StandardCharsets.UTF_8);
	try (Arena a = Arena.ofConfined()) {
	    MemorySegment ids = a.allocate(Math.max(id.length, 1));
	    MemorySegment.copy(id, 0, ids, JAVA_BYTE, 0, id.length);
	    MemorySegment off = a.allocate(JAVA_LONG, 2);
	    off.setAtIndex(JAVA_LONG, 0, 0L);
	    off.setAtIndex(JAVA_LONG, 1, id.length);
	    MemorySegment dup = a.allocate(1);
	    ctx.check((int) SwGpu.DEDUP_PROBE.invokeExact(ctx.handle, ids, off, 1, dup));
	    if (dup.get(JAVA_BYTE, 0) == 1) {
		return true; // shown before by this source
	    }
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable t) {
	    throw new SiteWhereException(t);
	}
	return stock.isDuplicate(request); // new to this source (or not decidable here): the store decides
    }

    public void setDeviceOrdinal(int deviceOrdinal) {
	this.deviceOrdinal = deviceOrdinal;
    }
}
