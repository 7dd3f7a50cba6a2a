/*
 * Drop-in for ProtobufDeviceEventDecoder
 * (service-event-sources/src/main/java/com/sitewhere/sources/decoder/protobuf/ProtobufDeviceEventDecoder.java:62,75-304):
 * same SPI, same results; the wire walk runs on the GPU for a micro-batch of concurrent callers.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import java.nio.charset.StandardCharsets;
import java.util.ArrayList;
import java.util.Date;
import java.util.HashMap;
import java.util.List;
import java.util.Map;

import com.sitewhere.rest.model.device.event.request.DeviceAlertCreateRequest;
import com.sitewhere.rest.model.device.event.request.DeviceEventCreateRequest;
import com.sitewhere.rest.model.device.event.request.DeviceLocationCreateRequest;
import com.sitewhere.rest.model.device.event.request.DeviceMeasurementCreateRequest;
import com.sitewhere.server.lifecycle.TenantEngineLifecycleComponent;
import com.sitewhere.sources.DecodedDeviceRequest;
import com.sitewhere.sources.decoder.protobuf.ProtobufDeviceEventDecoder;
import com.sitewhere.sources.spi.EventDecodeException;
import com.sitewhere.sources.spi.IDecodedDeviceRequest;
import com.sitewhere.sources.spi.IDeviceEventDecoder;
import com.sitewhere.spi.SiteWhereException;
import com.sitewhere.spi.device.event.AlertLevel;
import com.sitewhere.spi.device.event.AlertSource;
import com.sitewhere.spi.server.lifecycle.ILifecycleProgressMonitor;
import com.sitewhere.spi.server.lifecycle.LifecycleComponentType;

public class GpuProtobufDeviceEventDecoder extends TenantEngineLifecycleComponent implements IDeviceEventDecoder<byte[]> {

    static final int KIND_MEASUREMENT = 0, KIND_LOCATION = 1, KIND_ALERT = 2;
    static final int F_UPDATE_STATE = 1, F_HAS_DATE = 2, F_HAS_ORIGINATOR = 4, F_HAS_METADATA = 8;

    private int deviceOrdinal = 0;
    private int maxBatch = 65536;
    private long lingerMicros = 200;
    private SwGpu.Context ctx;
    private GpuBatcher batcher;
    /** non-event commands (registration, acknowledge, streams) keep going through the stock decoder */
    private final ProtobufDeviceEventDecoder stock = new ProtobufDeviceEventDecoder();

    public GpuProtobufDeviceEventDecoder() {
	super(LifecycleComponentType.DeviceEventDecoder);
    }

    @Override
    public void start(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	ctx = new SwGpu.Context(deviceOrdinal);
	batcher = new GpuBatcher(ctx, SwGpu.FMT_PROTOBUF, maxBatch, lingerMicros);
    }

    @Override
    public void stop(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	batcher.close();
	ctx.close();
    }

    @Override
    public List<IDecodedDeviceRequest<?>> decode(byte[] payload, Map<String, Object> metadata) throws EventDecodeException {
	GpuBatcher.Outcome out;
	try {
	    out = batcher.submit(payload).get();
	} catch (Exception e) {
	    throw new EventDecodeException("GPU decode failed.", e);
	}
	switch (out.status()) {
	case SwGpu.ST_DECODE_EXCEPTION:
	    throw new EventDecodeException("Unable to decode protobuf message."); // as :301-303
	case SwGpu.ST_RUNTIME_EXCEPTION:
	    throw new NullPointerException(); // parseDelimitedFrom returned null (:79-81)
	case SwGpu.ST_PASSTHROUGH:
	    return stock.decode(payload, metadata);
	default:
	    break;
	}
	List<IDecodedDeviceRequest<?>> results = new ArrayList<>();
	for (GpuBatcher.Row r : out.rows()) {
	    int kind = (int) (r.meta() & 0xff), flags = (int) ((r.meta() >> 8) & 0xffff);
	    DeviceEventCreateRequest request;
	    if (kind == KIND_LOCATION) {
		DeviceLocationCreateRequest loc = new DeviceLocationCreateRequest();
		loc.setLatitude(r.v0());
		loc.setLongitude(r.v1());
		loc.setElevation(r.v2());
		request = loc;
	    } else if (kind == KIND_ALERT) {
		DeviceAlertCreateRequest alert = new DeviceAlertCreateRequest();
		alert.setType(span(payload, r.nameSpan()));
		alert.setMessage(span(payload, r.messageSpan()));
		alert.setLevel(AlertLevel.Info);
		alert.setSource(AlertSource.Device);
		request = alert;
	    } else {
		DeviceMeasurementCreateRequest mx = new DeviceMeasurementCreateRequest();
		mx.setName(span(payload, r.nameSpan()));
		mx.setValue(r.v0());
		request = mx;
	    }
	    if ((flags & F_UPDATE_STATE) != 0) {
		request.setUpdateState(true);
	    }
	    request.setMetadata((flags & F_HAS_METADATA) != 0 ? PbMetadata.read(payload, r.metadataSpan()) : new HashMap<>());
	    request.setEventDate((flags & F_HAS_DATE) != 0 ? new Date(r.eventDate()) : new Date()); // :143-147
	    DecodedDeviceRequest<Object> decoded = new DecodedDeviceRequest<>();
	    if ((flags & F_HAS_ORIGINATOR) != 0) {
		decoded.setOriginator(span(payload, r.originatorSpan()));
	    }
	    decoded.setDeviceToken(span(payload, r.tokenSpan()));
	    decoded.setRequest(request);
	    results.add(decoded);
	}
	return results;
    }

    static String span(byte[] payload, long s) {
	return new String(payload, (int) (s & 0xffffffffL), (int) (s >>> 32), StandardCharsets.UTF_8);
    }

    public void setDeviceOrdinal(int deviceOrdinal) {
	this.deviceOrdinal = deviceOrdinal;
    }

    public void setMaxBatch(int maxBatch) {
	this.maxBatch = maxBatch;
    }

    public void setLingerMicros(long lingerMicros) {
	this.lingerMicros = lingerMicros;
    }
}
