/*
 * Drop-in for JsonDeviceRequestDecoder
 * (service-event-sources/src/main/java/com/sitewhere/sources/decoder/json/JsonDeviceRequestDecoder.java:48-62 and
 * sitewhere-core/.../marshaling/JsonDeviceRequestMarshaler.java:57-160): same SPI, same results; the tokenizer,
 * the bean rules and the number / date conversions run on the GPU for a micro-batch of concurrent callers.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import java.nio.charset.StandardCharsets;
import java.util.ArrayList;
import java.util.Date;
import java.util.List;
import java.util.Map;

import com.sitewhere.rest.model.device.event.request.DeviceAlertCreateRequest;
import com.sitewhere.rest.model.device.event.request.DeviceEventCreateRequest;
import com.sitewhere.rest.model.device.event.request.DeviceLocationCreateRequest;
import com.sitewhere.rest.model.device.event.request.DeviceMeasurementCreateRequest;
import com.sitewhere.server.lifecycle.TenantEngineLifecycleComponent;
import com.sitewhere.sources.DecodedDeviceRequest;
import com.sitewhere.sources.decoder.json.JsonDeviceRequestDecoder;
import com.sitewhere.sources.spi.EventDecodeException;
import com.sitewhere.sources.spi.IDecodedDeviceRequest;
import com.sitewhere.sources.spi.IDeviceEventDecoder;
import com.sitewhere.spi.SiteWhereException;
import com.sitewhere.spi.device.event.AlertLevel;
import com.sitewhere.spi.device.event.AlertSource;
import com.sitewhere.spi.device.event.DeviceEventType;
import com.sitewhere.spi.server.lifecycle.ILifecycleProgressMonitor;
import com.sitewhere.spi.server.lifecycle.LifecycleComponentType;

public class GpuJsonDeviceRequestDecoder extends TenantEngineLifecycleComponent implements IDeviceEventDecoder<byte[]> {

    static final int KIND_MEASUREMENT = 0, KIND_LOCATION = 1, KIND_ALERT = 2;
    // include/swgpu.h SWGPU_F_*
    static final int F_UPDATE_STATE = 1 << 0, F_HAS_DATE = 1 << 1, F_HAS_ORIGINATOR = 1 << 2, F_HAS_METADATA = 1 << 3,
	    F_HAS_ALT_ID = 1 << 4, F_NULL_TOKEN = 1 << 5, F_HAS_V0 = 1 << 6, F_HAS_V1 = 1 << 7, F_HAS_V2 = 1 << 8,
	    F_STR_ESCAPED = 1 << 9, F_HAS_NAME = 1 << 10, F_HAS_MESSAGE = 1 << 11, F_NAME_COERCED = 1 << 12,
	    F_NULL_REQUEST = 1 << 13, F_NULL_LEVEL = 1 << 14, F_NULL_SOURCE = 1 << 15;

    private int deviceOrdinal = 0;
    private int maxBatch = 65536;
    private long lingerMicros = 200;
    private SwGpu.Context ctx;
    private GpuBatcher batcher;
    /**
     * Everything the kernel hands back (SWGPU_ST_PASSTHROUGH: non-event request types, escaped key strings, exotic
     * encodings or number / date spellings — the list is in DESIGN.md) takes the stock Jackson path, which then
     * produces the reference's own result or exception for it.
     */
    private final JsonDeviceRequestDecoder stock = new JsonDeviceRequestDecoder();

    public GpuJsonDeviceRequestDecoder() {
	super(LifecycleComponentType.DeviceEventDecoder);
    }

    @Override
    public void start(ILifecycleProgressMonitor monitor) throws SiteWhereException {
	ctx = new SwGpu.Context(deviceOrdinal);
	batcher = new GpuBatcher(ctx, SwGpu.FMT_JSON, maxBatch, lingerMicros);
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
	if (out.status() != SwGpu.ST_OK) {
	    // DECODE_EXCEPTION / RUNTIME_EXCEPTION: the stock decoder throws the reference's own exception, with its own
	    // message and cause, for exactly these payloads (the kernel's verdict is pinned against it by the parity
	    // tests); PASSTHROUGH: it decodes what the kernel does not.
	    return stock.decode(payload, metadata);
	}
	List<IDecodedDeviceRequest<?>> results = new ArrayList<>();
	for (GpuBatcher.Row r : out.rows()) {
	    int kind = (int) (r.meta() & 0xff), flags = (int) ((r.meta() >> 8) & 0xffff);
	    if ((flags & (F_STR_ESCAPED | F_NAME_COERCED)) != 0) {
		// some string value holds backslash escapes, or a string property was given as a number / boolean
		// token: spans are raw payload bytes, so let Jackson materialise those strings
		return stock.decode(payload, metadata);
	    }
	    DeviceEventCreateRequest request;
	    if (kind == KIND_LOCATION) {
		DeviceLocationCreateRequest loc = new DeviceLocationCreateRequest();
		if ((flags & F_HAS_V0) != 0) {
		    loc.setLatitude(r.v0());
		}
		if ((flags & F_HAS_V1) != 0) {
		    loc.setLongitude(r.v1());
		}
		if ((flags & F_HAS_V2) != 0) {
		    loc.setElevation(r.v2());
		}
		request = loc;
	    } else if (kind == KIND_ALERT) {
		DeviceAlertCreateRequest alert = new DeviceAlertCreateRequest();
		if ((flags & F_HAS_NAME) != 0) {
		    alert.setType(span(payload, r.nameSpan()));
		}
		if ((flags & F_HAS_MESSAGE) != 0) {
		    alert.setMessage(span(payload, r.messageSpan()));
		}
		// meta carries the effective enum ordinals (the bean's initialisers when the property is absent);
		// an explicit JSON null clears the field
		alert.setLevel((flags & F_NULL_LEVEL) != 0 ? null : AlertLevel.values()[(int) ((r.meta() >> 24) & 3)]);
		alert.setSource((flags & F_NULL_SOURCE) != 0 ? null : AlertSource.values()[(int) ((r.meta() >> 26) & 1)]);
		request = alert;
	    } else {
		DeviceMeasurementCreateRequest mx = new DeviceMeasurementCreateRequest();
		if ((flags & F_HAS_NAME) != 0) {
		    mx.setName(span(payload, r.nameSpan()));
		}
		if ((flags & F_HAS_V0) != 0) {
		    mx.setValue(r.v0());
		}
		request = mx;
	    }
	    if (((r.meta() >> 27) & 1) != 0) {
		request.setEventType(DeviceEventType.values()[(int) ((r.meta() >> 28) & 7)]);
	    }
	    if ((flags & F_HAS_ALT_ID) != 0) {
		request.setAlternateId(span(payload, r.alternateIdSpan()));
	    }
	    if ((flags & F_UPDATE_STATE) != 0) {
		request.setUpdateState(true);
	    }
	    if ((flags & F_HAS_METADATA) != 0) {
		// the span is the text of the metadata object, validated by the library as a flat object of scalar
		// values: Jackson reads it into the Map<String, String> exactly as the reference's treeToValue does
		long ms = r.metadataSpan();
		try {
		    request.setMetadata(METADATA_MAPPER.readValue(payload, (int) (ms & 0xffffffffL), (int) (ms >>> 32),
			    METADATA_TYPE));
		} catch (java.io.IOException e) {
		    throw new EventDecodeException(e);
		}
	    }
	    if ((flags & F_HAS_DATE) != 0) {
		request.setEventDate(new Date(r.eventDate()));
	    }
	    DecodedDeviceRequest<Object> decoded = new DecodedDeviceRequest<>();
	    if ((flags & F_HAS_ORIGINATOR) != 0) {
		decoded.setOriginator(span(payload, r.originatorSpan()));
	    }
	    if ((flags & F_NULL_TOKEN) == 0) {
		decoded.setDeviceToken(span(payload, r.tokenSpan()));
	    }
	    decoded.setRequest((flags & F_NULL_REQUEST) != 0 ? null : request);
	    results.add(decoded);
	}
	return results;
    }

    private static final com.fasterxml.jackson.databind.ObjectMapper METADATA_MAPPER = new com.fasterxml.jackson.databind.ObjectMapper();
    private static final com.fasterxml.jackson.databind.JavaType METADATA_TYPE = METADATA_MAPPER.getTypeFactory()
	    .constructMapType(java.util.HashMap.class, String.class, String.class);

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
