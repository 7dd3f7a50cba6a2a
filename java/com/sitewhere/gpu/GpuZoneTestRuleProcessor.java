/*
 * Drop-in for ZoneTestRuleProcessor
 * (service-rule-processing/src/main/java/com/sitewhere/rules/processors/geospatial/ZoneTestRuleProcessor.java:33-93):
 * same `zoneTests` property, so the existing <zone-test-processor> configuration binds unchanged
 * (RuleProcessingParser.java:103-152).  Locations are batched; one swgpu_zone_test call answers a batch, and
 * alerts leave through the same getDeviceEventManagement().addDeviceAlerts call, in the same order.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.util.ArrayList;
import java.util.Date;
import java.util.List;

import com.sitewhere.rest.model.device.event.request.DeviceAlertCreateRequest;
import com.sitewhere.rules.processors.geospatial.RuleProcessor;
import com.sitewhere.rules.processors.geospatial.ZoneTest;
import com.sitewhere.spi.SiteWhereException;
import com.sitewhere.spi.area.IZone;
import com.sitewhere.spi.common.ILocation;
import com.sitewhere.spi.device.event.IDeviceEventContext;
import com.sitewhere.spi.device.event.IDeviceLocation;
import com.sitewhere.spi.device.event.ZoneContainment;
import com.sitewhere.spi.server.lifecycle.ILifecycleProgressMonitor;

public class GpuZoneTestRuleProcessor extends RuleProcessor {

    private List<ZoneTest> zoneTests = new ArrayList<ZoneTest>();
    private int deviceOrdinal = 0;
    private int maxBatch = 65536;
    /** a location waits at most this long for company before the pending batch is tested */
    private long lingerMicros = 2000;
    private SwGpu.Context ctx;
    private final List<IDeviceLocation> pending = new ArrayList<>();
    private java.util.concurrent.ScheduledExecutorService lingerTimer;

    @Override
    public void start(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	super.start(monitor);
	ctx = new SwGpu.Context(deviceOrdinal);
	uploadZones();
	// The stock host (KafkaRuleProcessorHost.java:144-153,221-224) calls onLocation once per event from its
	// thread pool and nothing else, so the flush cannot depend on the host: a timer tests whatever is
	// pending every lingerMicros (alerts are delayed by at most that), and a full batch flushes at once.
	lingerTimer = java.util.concurrent.Executors.newSingleThreadScheduledExecutor(r -> {
	    Thread t = new Thread(r, "gpu-zone-test-linger");
	    t.setDaemon(true);
	    return t;
	});
	lingerTimer.scheduleWithFixedDelay(() -> {
	    try {
		flush();
	    } catch (Throwable e) {
		getLogger().error("GPU zone test flush failed.", e);
	    }
	}, lingerMicros, lingerMicros, java.util.concurrent.TimeUnit.MICROSECONDS);
    }

    @Override
    public void stop(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	if (lingerTimer != null) {
	    lingerTimer.shutdownNow();
	}
	flush();
	ctx.close();
	super.stop(monitor);
    }

    /** Flatten every test's zone bounds (as GeoUtils.createPolygonForLocations reads them) and upload. */
    public void uploadZones() throws SiteWhereException {
	try (Arena a = Arena.ofConfined()) {
	    List<Double> lon = new ArrayList<>(), lat = new ArrayList<>();
	    MemorySegment off = a.allocate(JAVA_INT, zoneTests.size() + 1L);
	    MemorySegment cond = a.allocate(Math.max(zoneTests.size(), 1));
	    int t = 0;
	    for (ZoneTest test : zoneTests) {
		IZone zone = getDeviceManagement().getZoneByToken(test.getZoneToken());
		if (zone == null) {
		    throw new SiteWhereException("Invalid zone token in " + getClass().getName() + ": " + test.getZoneToken());
		}
		off.setAtIndex(JAVA_INT, t, lon.size());
		for (ILocation loc : zone.getBounds()) {
		    lon.add(loc.getLongitude());
		    lat.add(loc.getLatitude());
		}
		cond.set(JAVA_BYTE, t, (byte) (test.getCondition() == ZoneContainment.Inside ? 0 : 1));
		t++;
	    }
	    off.setAtIndex(JAVA_INT, t, lon.size());
	    MemorySegment lo = a.allocate(JAVA_DOUBLE, Math.max(lon.size(), 1)), la = a.allocate(JAVA_DOUBLE, Math.max(lat.size(), 1));
	    for (int i = 0; i < lon.size(); i++) {
		lo.setAtIndex(JAVA_DOUBLE, i, lon.get(i));
		la.setAtIndex(JAVA_DOUBLE, i, lat.get(i));
	    }
	    ctx.check((int) SwGpu.SET_ZONES.invokeExact(ctx.handle, zoneTests.size(), off, lo, la, cond));
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable e) {
	    throw new SiteWhereException(e);
	}
    }

    @Override
    public synchronized void onLocation(IDeviceEventContext context, IDeviceLocation location) throws SiteWhereException {
	pending.add(location);
	if (pending.size() >= maxBatch) {
	    flush();
	}
    }

    /** Test all pending locations at once: called by the linger timer, and by onLocation when a batch is full. */
    public synchronized void flush() throws SiteWhereException {
	int n = pending.size();
	if (n == 0) {
	    return;
	}
	try (Arena a = Arena.ofConfined()) {
	    MemorySegment lon = a.allocate(JAVA_DOUBLE, n), lat = a.allocate(JAVA_DOUBLE, n);
	    for (int i = 0; i < n; i++) {
		lon.setAtIndex(JAVA_DOUBLE, i, pending.get(i).getLongitude());
		lat.setAtIndex(JAVA_DOUBLE, i, pending.get(i).getLatitude());
	    }
	    long cap = Math.max(1024, (long) n * Math.max(1, zoneTests.size()));
	    MemorySegment pairs = a.allocate(JAVA_LONG, cap), np = a.allocate(JAVA_LONG);
	    ctx.check((int) SwGpu.ZONE_TEST.invokeExact(ctx.handle, lon, lat, MemorySegment.NULL, n, cap, pairs, np));
	    long count = np.get(JAVA_LONG, 0);
	    for (long k = 0; k < count; k++) { // (location << 32 | test), in the reference's firing order
		long p = pairs.getAtIndex(JAVA_LONG, k);
		IDeviceLocation location = pending.get((int) (p >>> 32));
		ZoneTest test = zoneTests.get((int) (p & 0xffffffffL));
		DeviceAlertCreateRequest alert = new DeviceAlertCreateRequest();
		alert.setType(test.getAlertType());
		alert.setLevel(test.getAlertLevel());
		alert.setMessage(test.getAlertMessage());
		alert.setUpdateState(false);
		alert.setEventDate(new Date());
		getDeviceEventManagement().addDeviceAlerts(location.getDeviceAssignmentId(), alert);
	    }
	} catch (SiteWhereException e) {
	    throw e;
	} catch (Throwable e) {
	    throw new SiteWhereException(e);
	} finally {
	    pending.clear();
	}
    }

    public List<ZoneTest> getZoneTests() {
	return zoneTests;
    }

    public void setZoneTests(List<ZoneTest> zoneTests) {
	this.zoneTests = zoneTests;
    }

    public void setDeviceOrdinal(int deviceOrdinal) {
	this.deviceOrdinal = deviceOrdinal;
    }

    public void setLingerMicros(long lingerMicros) {
	this.lingerMicros = lingerMicros;
    }

    public void setMaxBatch(int maxBatch) {
	this.maxBatch = maxBatch;
    }
}
