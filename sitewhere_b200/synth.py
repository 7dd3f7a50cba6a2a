"""Synthetic inputs for the hot path: device tokens, protobuf and JSON payload
batches in the wire shapes the reference itself emits, and zone tables.

Wire shapes follow the reference's own encoders:
  * protobuf: ``ProtobufDeviceEventEncoder`` (service-event-sources/src/main/java/com/sitewhere/
    sources/decoder/protobuf/ProtobufDeviceEventEncoder.java:60-176) — a delimited
    ``SiteWhere.Header{command}`` followed by a delimited ``Model.*`` message with hardwareId,
    eventDate and the event fields, schema sitewhere-communication/src/main/proto/sitewhere.proto.
  * JSON: the ``DeviceRequest`` envelope (sitewhere-core-api/.../rest/model/device/communication/
    DeviceRequest.java:15-88) as a default Jackson ``ObjectMapper`` writes it, with ``eventDate`` in
    the ``JsonDateSerializer`` form ``yyyy-MM-dd'T'HH:mm:ss.SSS'Z'``.

Everything is deterministic in ``seed`` (numpy PCG64) and vectorised so that the 10 M-record
benchmark batch is built in seconds.  All generators return ``(bytes: uint8[total], rec_off:
uint64[n+1])`` plus the ground-truth columns used to build them.
"""
from __future__ import annotations

import math
import struct

import numpy as np

SEED_BASE = 0x5157_2026
EVENT_DATE_BASE = 1_527_854_400_000  # 2018-06-01T12:00:00Z, ms

MEASUREMENT_NAMES = [
    "fuel.level", "engine.temp", "oil.psi", "rpm.avg", "speed.kmh", "battery.v", "cabin.temp",
    "load.kg", "tire.psi.fl", "tire.psi.fr", "coolant.t", "humidity", "vibration", "pressure.kpa",
    "flow.lpm", "signal.dbm",
]
ALERT_TYPES = [
    "engine.overheat", "fuel.low", "geofence.exit", "battery.low", "tamper", "speeding",
    "maintenance", "offline.warn",
]


# ------------------------------------------------------------------------------------------ tokens
def token_bytes(dev: np.ndarray) -> np.ndarray:
    """Device tokens ``%05d-SYNTH-%09d`` (21 chars, the length of the reference fixtures' tokens such
    as ``20544-OPENHAB-7313064``) as a uint8[n, 21] matrix."""
    dev = np.asarray(dev, dtype=np.int64)
    n = dev.shape[0]
    out = np.empty((n, 21), dtype=np.uint8)
    pre = (dev * 7919 + 104729) % 100000
    for k in range(5):
        out[:, 4 - k] = 48 + (pre // (10 ** k)) % 10
    out[:, 5:12] = np.frombuffer(b"-SYNTH-", dtype=np.uint8)
    for k in range(9):
        out[:, 20 - k] = 48 + (dev // (10 ** k)) % 10
    return out


def token_str(d: int) -> str:
    return bytes(token_bytes(np.array([d]))[0]).decode()


def murmur2_rows(tok: np.ndarray) -> np.ndarray:
    """Kafka Utils.murmur2 of each row of a uint8[n, L] matrix (vectorised; same L for all rows)."""
    tok = np.ascontiguousarray(tok, dtype=np.uint8)
    n, ln = tok.shape
    m = np.uint32(0x5BD1E995)
    h = np.full(n, 0x9747B28C ^ ln, dtype=np.uint32)
    t = tok.astype(np.uint32)
    with np.errstate(over="ignore"):
        for i in range(ln // 4):
            k = t[:, 4 * i] | (t[:, 4 * i + 1] << 8) | (t[:, 4 * i + 2] << 16) | (t[:, 4 * i + 3] << 24)
            k = k * m
            k ^= k >> np.uint32(24)
            k = k * m
            h = h * m
            h ^= k
        base, rem = ln & ~3, ln & 3
        if rem:
            tail = t[:, base].copy()
            if rem > 1:
                tail |= t[:, base + 1] << 8
            if rem > 2:
                tail |= t[:, base + 2] << 16
            h ^= tail
            h = h * m
        h ^= h >> np.uint32(13)
        h = h * m
        h ^= h >> np.uint32(15)
    return h


def devices_of_shard(n_devices: int, rank: int, world: int) -> np.ndarray:
    """Device indices whose token Kafka's DefaultPartitioner routes to partition ``rank`` of ``world``:
    toPositive(murmur2(token)) % world (sitewhere-microservice/.../kafka/MicroserviceKafkaProducer.java:89-92
    sends key = deviceToken without a partition)."""
    dev = np.arange(n_devices, dtype=np.int64)
    if world <= 1:
        return dev
    h = murmur2_rows(token_bytes(dev))
    return dev[((h & np.uint32(0x7FFFFFFF)) % np.uint32(world)) == rank]


# ------------------------------------------------------------------------------------------ zones
def star_zones(grid: int, nvert: int = 64, lon0=-85.0, lon1=-84.0, lat0=34.0, lat1=35.0,
               ratio: float = 0.5, fill: float = 0.45, seed: int = SEED_BASE + 1, cond_mode="inside"):
    """``grid*grid`` simple concave star polygons (``nvert`` vertices, inner/outer radius ``ratio``),
    centres on a regular grid.  Returns (vtx_off uint32[T+1], lon f64[], lat f64[], cond uint8[T])
    with OPEN rings, as ``IZone.getBounds()`` holds them."""
    rng = np.random.default_rng(seed)
    T = grid * grid
    sx = (lon1 - lon0) / grid
    sy = (lat1 - lat0) / grid
    cx = lon0 + (np.arange(T) % grid + 0.5) * sx
    cy = lat0 + (np.arange(T) // grid + 0.5) * sy
    phase = rng.uniform(0, 2 * math.pi, size=T)
    k = np.arange(nvert)
    ang = phase[:, None] + 2 * math.pi * k[None, :] / nvert
    rad = np.where(k % 2 == 0, 1.0, ratio)[None, :]
    lon = cx[:, None] + fill * sx * rad * np.cos(ang)
    lat = cy[:, None] + fill * sy * rad * np.sin(ang)
    vtx_off = (np.arange(T + 1) * nvert).astype(np.uint32)
    if cond_mode == "inside":
        cond = np.zeros(T, dtype=np.uint8)
    elif cond_mode == "alternate":
        cond = (np.arange(T) % 2).astype(np.uint8)
    else:
        cond = np.ones(T, dtype=np.uint8)
    return vtx_off, np.ascontiguousarray(lon.reshape(-1)), np.ascontiguousarray(lat.reshape(-1)), cond


def rect_zones(grid: int = 4, lon0=-85.0, lon1=-84.0, lat0=34.0, lat1=35.0, fill=0.8):
    """C1: ``grid*grid`` axis-aligned rectangles (JTS ``isRectangle`` path), conditions alternating."""
    T = grid * grid
    sx = (lon1 - lon0) / grid
    sy = (lat1 - lat0) / grid
    lon = np.empty((T, 4))
    lat = np.empty((T, 4))
    for t in range(T):
        x0 = lon0 + (t % grid + (1 - fill) / 2) * sx
        y0 = lat0 + (t // grid + (1 - fill) / 2) * sy
        x1, y1 = x0 + fill * sx, y0 + fill * sy
        lon[t] = [x0, x1, x1, x0]
        lat[t] = [y0, y0, y1, y1]
    vtx_off = (np.arange(T + 1) * 4).astype(np.uint32)
    cond = (np.arange(T) % 2).astype(np.uint8)
    return vtx_off, lon.reshape(-1).copy(), lat.reshape(-1).copy(), cond


def locations_for_zones(n: int, vtx_off, zlon, zlat, seed: int, inflate=0.10, boundary_rate=1e-4):
    """Uniform points in the zones' bounding box inflated by ``inflate``; a ``boundary_rate`` share
    are placed exactly on zone vertices, edge midpoints' coordinates, or envelope borders."""
    rng = np.random.default_rng(seed)
    x0, x1 = zlon.min(), zlon.max()
    y0, y1 = zlat.min(), zlat.max()
    dx, dy = (x1 - x0) * inflate / 2, (y1 - y0) * inflate / 2
    lon = rng.uniform(x0 - dx, x1 + dx, size=n)
    lat = rng.uniform(y0 - dy, y1 + dy, size=n)
    nb = int(n * boundary_rate)
    if nb:
        idx = rng.choice(n, size=nb, replace=False)
        v = rng.integers(0, zlon.shape[0], size=nb)
        mode = rng.integers(0, 3, size=nb)
        # 0: exactly a vertex; 1: same y as a vertex (horizontal ray through a vertex); 2: same x
        lon[idx] = np.where(mode != 1, zlon[v], lon[idx])
        lat[idx] = np.where(mode != 2, zlat[v], lat[idx])
    return lon, lat


# ------------------------------------------------------------------------------------------ protobuf
def _f64_bytes(a: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(a, dtype="<f8").view(np.uint8).reshape(-1, 8)


def _u64_bytes(a: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(a, dtype="<u8").view(np.uint8).reshape(-1, 8)


def pb_locations(n: int, n_devices: int, lon: np.ndarray, lat: np.ndarray, seed: int, date0=EVENT_DATE_BASE,
                 devices=None):
    """n ``SEND_DEVICE_LOCATION`` payloads (63 B each).  Returns bytes, rec_off, truth dict.
    ``devices``: optional array of device indices to draw from (a shard's devices)."""
    rng = np.random.default_rng(seed)
    dev = rng.integers(0, n_devices, size=n) if devices is None else np.asarray(devices)[rng.integers(0, len(devices), size=n)]
    elev = np.round(rng.uniform(0, 500, size=n), 1)
    date = date0 + np.arange(n, dtype=np.int64)
    rec = np.empty((n, 63), dtype=np.uint8)
    rec[:, 0:3] = (2, 0x08, 3)          # Header: len 2, command = SEND_DEVICE_LOCATION
    rec[:, 3] = 59                      # DeviceLocation length
    rec[:, 4:6] = (0x0A, 21)            # hardwareId
    rec[:, 6:27] = token_bytes(dev)
    rec[:, 27] = 0x11                   # latitude
    rec[:, 28:36] = _f64_bytes(lat)
    rec[:, 36] = 0x19                   # longitude
    rec[:, 37:45] = _f64_bytes(lon)
    rec[:, 45] = 0x21                   # elevation
    rec[:, 46:54] = _f64_bytes(elev)
    rec[:, 54] = 0x29                   # eventDate fixed64
    rec[:, 55:63] = _u64_bytes(date.astype(np.uint64))
    rec_off = (np.arange(n + 1, dtype=np.uint64) * 63)
    return rec.reshape(-1), rec_off, dict(dev=dev, lat=lat, lon=lon, elev=elev, date=date)


def enriched_locations(n: int, n_assignments: int, lon: np.ndarray, lat: np.ndarray, seed: int, date0=EVENT_DATE_BASE):
    """n serialised ``GEnrichedEventPayload`` records holding a location event (159 B each), as the
    ``inbound-enriched-events`` topic carries them (device-event-model.proto:139-150, 171-176, 204-213, 288-296,
    321-324): context {deviceId, deviceTypeId, assignmentStatus = ACTIVE}, event.location {latitude, longitude,
    elevation, event {id, eventType = LOCATION, deviceId, deviceAssignmentId, eventDate, receivedDate}}.  Event
    dates must lie in [2^35, 2^42) ms (six-byte varints).  Returns bytes, rec_off, truth dict."""
    rng = np.random.default_rng(seed)
    asg = rng.integers(0, n_assignments, size=n).astype(np.uint64)
    elev = np.round(rng.uniform(0, 500, size=n), 1)
    date = (date0 + np.arange(n, dtype=np.int64)).astype(np.uint64)
    assert int(date.min()) >= 1 << 35 and int(date.max()) < 1 << 42
    ev_id = np.arange(n, dtype=np.uint64) + np.uint64(seed) * np.uint64(1 << 40)

    def guuid(rec, at, tag, msb, lsb):       # tag 0x12 (len 18): 09 msb(fixed64) 11 lsb(fixed64)
        rec[:, at:at + 2] = (tag, 18)
        rec[:, at + 2] = 0x09
        rec[:, at + 3:at + 11] = _u64_bytes(msb)
        rec[:, at + 11] = 0x11
        rec[:, at + 12:at + 20] = _u64_bytes(lsb)
        return at + 20

    def optd(rec, at, tag, v):                # GOptionalDouble{value}: tag 09 09 f64  (values are never 0.0 here)
        rec[:, at:at + 3] = (tag, 9, 0x09)
        rec[:, at + 3:at + 11] = _f64_bytes(v)
        return at + 11

    def varint6(rec, at, tag, v):
        rec[:, at] = tag
        for k in range(6):
            rec[:, at + 1 + k] = ((v >> np.uint64(7 * k)) & np.uint64(0x7f)).astype(np.uint8) | (0x80 if k < 5 else 0)
        return at + 7
    L = 159
    rec = np.empty((n, L), dtype=np.uint8)
    one = np.ones(n, dtype=np.uint64)
    rec[:, 0:2] = (0x0A, 42)                 # context
    at = guuid(rec, 2, 0x0A, asg * np.uint64(3) + one, asg)               # deviceId
    at = guuid(rec, at, 0x12, one * np.uint64(77), one * np.uint64(78))   # deviceTypeId
    rec[:, at:at + 2] = (0x30, 1)            # assignmentStatus = ACTIVE
    at += 2
    rec[:, at:at + 2] = (0x12, 113)          # event (GAnyDeviceEvent)
    rec[:, at + 2:at + 4] = (0x1A, 111)      # location
    at += 4
    lat_ = np.where(np.asarray(lat) == 0.0, 1e-9, lat)
    lon_ = np.where(np.asarray(lon) == 0.0, 1e-9, lon)
    at = optd(rec, at, 0x0A, lat_)
    at = optd(rec, at, 0x12, lon_)
    at = optd(rec, at, 0x1A, np.where(elev == 0.0, 0.5, elev))
    rec[:, at:at + 2] = (0x22, 76)           # GDeviceEvent
    at += 2
    at = guuid(rec, at, 0x0A, ev_id, ev_id ^ np.uint64(0x5555))           # id
    rec[:, at:at + 2] = (0x18, 1)            # eventType = EVENT_TYPE_LOCATION
    at += 2
    at = guuid(rec, at, 0x22, asg * np.uint64(3) + one, asg)              # deviceId
    at = guuid(rec, at, 0x2A, asg + np.uint64(1 << 32), asg * np.uint64(0x9E3779B1) + one)   # deviceAssignmentId
    at = varint6(rec, at, 0x48, date)        # eventDate
    at = varint6(rec, at, 0x50, date + np.uint64(5))                      # receivedDate
    assert at == L, at
    rec_off = (np.arange(n + 1, dtype=np.uint64) * L)
    return rec.reshape(-1), rec_off, dict(assignment=asg, lat=lat_, lon=lon_, date=date.astype(np.int64), event_id=ev_id)


def _ragged_from_groups(n, group_id, templates):
    """Assemble n variable-length records where record i is built from templates[group_id[i]]
    (a function (idx) -> uint8[len(idx), L_g])."""
    lens = np.array([t[0] for t in templates], dtype=np.int64)[group_id]
    rec_off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=rec_off[1:].view(np.int64))
    out = np.empty(int(rec_off[-1]), dtype=np.uint8)
    for g, (L, fn) in enumerate(templates):
        idx = np.nonzero(group_id == g)[0]
        if idx.size == 0:
            continue
        block = fn(idx)
        pos = rec_off[idx].astype(np.int64)[:, None] + np.arange(L)[None, :]
        out[pos.reshape(-1)] = block.reshape(-1)
    return out, rec_off


def pb_mixed(n: int, n_devices: int, seed: int, zones=None, mix=(0.70, 0.25, 0.05), date0=EVENT_DATE_BASE):
    """C3: mixed measurement / location / alert protobuf payloads, one Measurement per message as
    ProtobufDeviceEventEncoder.encodeDeviceMeasurements writes them (…Encoder.java:60-93)."""
    rng = np.random.default_rng(seed)
    dev = rng.integers(0, n_devices, size=n)
    kind = rng.choice(3, size=n, p=np.array(mix))            # 0 measurement 1 location 2 alert
    sub = np.where(kind == 0, rng.integers(0, len(MEASUREMENT_NAMES), size=n),
                   np.where(kind == 2, rng.integers(0, len(ALERT_TYPES), size=n), 0))
    value = np.round(rng.uniform(-1000, 1000, size=n), 3)
    if zones is not None:
        lon, lat = locations_for_zones(n, zones[0], zones[1], zones[2], seed + 17)
    else:
        lon = rng.uniform(-85, -84, size=n)
        lat = rng.uniform(34, 35, size=n)
    date = date0 + np.arange(n, dtype=np.int64)
    tok = token_bytes(dev)
    datb = _u64_bytes(date.astype(np.uint64))
    nM, nA = len(MEASUREMENT_NAMES), len(ALERT_TYPES)
    group = np.where(kind == 0, sub, np.where(kind == 1, nM, nM + 1 + sub))
    templates = []
    for name in MEASUREMENT_NAMES:
        nb = name.encode()
        mlen = 2 + len(nb) + 9
        body = 23 + 2 + mlen + 9
        L = 3 + 1 + body

        def fn(idx, nb=nb, mlen=mlen, body=body, L=L):
            r = np.empty((idx.size, L), dtype=np.uint8)
            r[:, 0:3] = (2, 0x08, 5)
            r[:, 3] = body
            r[:, 4:6] = (0x0A, 21)
            r[:, 6:27] = tok[idx]
            r[:, 27:29] = (0x12, mlen)
            r[:, 29:31] = (0x0A, len(nb))
            r[:, 31:31 + len(nb)] = np.frombuffer(nb, dtype=np.uint8)
            p = 31 + len(nb)
            r[:, p] = 0x11
            r[:, p + 1:p + 9] = _f64_bytes(value[idx])
            r[:, p + 9] = 0x19
            r[:, p + 10:p + 18] = datb[idx]
            return r
        templates.append((L, fn))

    def fn_loc(idx):
        r = np.empty((idx.size, 63), dtype=np.uint8)
        r[:, 0:3] = (2, 0x08, 3)
        r[:, 3] = 59
        r[:, 4:6] = (0x0A, 21)
        r[:, 6:27] = tok[idx]
        r[:, 27] = 0x11
        r[:, 28:36] = _f64_bytes(lat[idx])
        r[:, 36] = 0x19
        r[:, 37:45] = _f64_bytes(lon[idx])
        r[:, 45] = 0x21
        r[:, 46:54] = 0
        r[:, 54] = 0x29
        r[:, 55:63] = datb[idx]
        return r
    templates.append((63, fn_loc))
    for at in ALERT_TYPES:
        ab = at.encode()
        msg = ("alert raised: " + at).encode()
        body = 23 + 2 + len(ab) + 2 + len(msg) + 9
        L = 3 + 1 + body

        def fn(idx, ab=ab, msg=msg, body=body, L=L):
            r = np.empty((idx.size, L), dtype=np.uint8)
            r[:, 0:3] = (2, 0x08, 4)
            r[:, 3] = body
            r[:, 4:6] = (0x0A, 21)
            r[:, 6:27] = tok[idx]
            p = 27
            r[:, p:p + 2] = (0x12, len(ab))
            r[:, p + 2:p + 2 + len(ab)] = np.frombuffer(ab, dtype=np.uint8)
            p += 2 + len(ab)
            r[:, p:p + 2] = (0x1A, len(msg))
            r[:, p + 2:p + 2 + len(msg)] = np.frombuffer(msg, dtype=np.uint8)
            p += 2 + len(msg)
            r[:, p] = 0x21
            r[:, p + 1:p + 9] = datb[idx]
            return r
        templates.append((L, fn))
    out, rec_off = _ragged_from_groups(n, group, templates)
    return out, rec_off, dict(dev=dev, kind=kind, sub=sub, value=value, lon=lon, lat=lat, date=date)


# ------------------------------------------------------------------------------------------ JSON
def iso_date(ms: int) -> str:
    """JsonDateSerializer.serialize: ISODateTimeFormat.dateTime() in UTC
    (sitewhere-core-api/.../rest/model/datatype/JsonDateSerializer.java:48-51)."""
    days, rem = divmod(ms, 86_400_000)
    z = days + 719468
    era = z // 146097
    doe = z - era * 146097
    yoe = (doe - doe // 1460 + doe // 36524 - doe // 146096) // 365
    y = yoe + era * 400
    doy = doe - (365 * yoe + yoe // 4 - yoe // 100)
    mp = (5 * doy + 2) // 153
    d = doy - (153 * mp + 2) // 5 + 1
    m = mp + 3 if mp < 10 else mp - 9
    if m <= 2:
        y += 1
    hh, rem = divmod(rem, 3_600_000)
    mm, rem = divmod(rem, 60_000)
    ss, ms_ = divmod(rem, 1000)
    return f"{y:04d}-{m:02d}-{d:02d}T{hh:02d}:{mm:02d}:{ss:02d}.{ms_:03d}Z"


def java_double_str(v: float) -> str:
    """Double.toString for the magnitudes used here (1e-3 <= |v| < 1e7 prints as plain decimal;
    shortest repr that round-trips, always with a fraction digit)."""
    if v == 0:
        return "-0.0" if math.copysign(1, v) < 0 else "0.0"
    a = abs(v)
    if 1e-3 <= a < 1e7:
        s = repr(float(v))
        if "e" in s or "E" in s:
            s = f"{v:.17g}"
        if "." not in s:
            s += ".0"
        return s
    # computerised scientific notation d.dddE[-]n
    m, e = f"{v:.17e}".split("e")
    r = repr(float(v))
    if "e" in r:
        m, e = r.split("e")
    else:
        m, e = f"{float(r):e}".split("e")
        # recover shortest digits
        digits = r.replace("-", "").replace(".", "").lstrip("0").rstrip("0") or "0"
        m = ("-" if v < 0 else "") + digits[0] + "." + (digits[1:] or "0")
    if "." not in m:
        m += ".0"
    return f"{m}E{int(e)}"


def json_measurement(token: str, name: str, value_text: str, date_ms: int) -> bytes:
    """The reference's own shape: DeviceRequest{deviceToken, originator, type, request} with a
    DeviceMeasurementCreateRequest serialised by a default ObjectMapper (field order = getter
    order of the POJOs; @JsonInclude(NON_NULL) drops null alternateId/metadata)."""
    return (
        '{"deviceToken":"%s","originator":null,"type":"DeviceMeasurement","request":'
        '{"eventType":"Measurement","eventDate":"%s","updateState":false,"name":"%s","value":%s}}'
        % (token, iso_date(date_ms), name, value_text)
    ).encode()


def json_location(token: str, lat_text: str, lon_text: str, elev_text: str, date_ms: int,
                  update_state=False, metadata=None) -> bytes:
    md = ""
    if metadata:
        md = ',"metadata":{' + ",".join('"%s":"%s"' % kv for kv in metadata.items()) + "}"
    return (
        '{"deviceToken":"%s","originator":null,"type":"DeviceLocation","request":'
        '{"eventType":"Location","eventDate":"%s","updateState":%s%s,"latitude":%s,"longitude":%s,"elevation":%s}}'
        % (token, iso_date(date_ms), "true" if update_state else "false", md, lat_text, lon_text, elev_text)
    ).encode()


def json_alert(token: str, atype: str, message: str, date_ms: int, level="Info", source="Device") -> bytes:
    return (
        '{"deviceToken":"%s","originator":null,"type":"DeviceAlert","request":'
        '{"eventType":"Alert","eventDate":"%s","updateState":false,"source":"%s","level":"%s","type":"%s","message":"%s"}}'
        % (token, iso_date(date_ms), source, level, atype, message)
    ).encode()


def pack_records(recs) -> tuple[np.ndarray, np.ndarray]:
    """list[bytes] -> (uint8[total], uint64[n+1])."""
    lens = np.fromiter((len(r) for r in recs), dtype=np.int64, count=len(recs))
    rec_off = np.zeros(len(recs) + 1, dtype=np.uint64)
    np.cumsum(lens, out=rec_off[1:].view(np.int64))
    data = np.frombuffer(b"".join(recs), dtype=np.uint8).copy() if recs else np.zeros(0, np.uint8)
    return data, rec_off


def json_measurements(n: int, n_devices: int, seed: int, date0=EVENT_DATE_BASE):
    """C1: n JSON DeviceMeasurement requests; 1 % values with 17 significant digits, 0.1 % in
    exponent form."""
    rng = np.random.default_rng(seed)
    dev = rng.integers(0, n_devices, size=n)
    nm = rng.integers(0, len(MEASUREMENT_NAMES), size=n)
    val = rng.uniform(-1000, 1000, size=n)
    style = rng.random(size=n)
    toks = token_bytes(dev)
    recs = []
    texts = []
    for i in range(n):
        v = float(val[i])
        if style[i] < 0.001:
            t = "%.6e" % v
        elif style[i] < 0.011:
            t = "%.17g" % v
            if "." not in t and "e" not in t:
                t += ".0"
        else:
            t = java_double_str(round(v, 3))
        texts.append(t)
        recs.append(json_measurement(bytes(toks[i]).decode(), MEASUREMENT_NAMES[nm[i]], t, int(date0 + i)))
    data, rec_off = pack_records(recs)
    return data, rec_off, dict(dev=dev, name=nm, value_text=texts, date=date0 + np.arange(n, dtype=np.int64))


# ------------------------------------------------------------------------------------------ scalar pb encoder
def varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def pb_field_len(field: int, payload: bytes) -> bytes:
    return varint((field << 3) | 2) + varint(len(payload)) + payload


def pb_field_f64(field: int, v: float) -> bytes:
    return varint((field << 3) | 1) + struct.pack("<d", v)


def pb_field_fixed64(field: int, v: int) -> bytes:
    return varint((field << 3) | 1) + struct.pack("<Q", v & ((1 << 64) - 1))


def pb_field_varint(field: int, v: int) -> bytes:
    return varint((field << 3) | 0) + varint(v)


def pb_delimited(msg: bytes) -> bytes:
    return varint(len(msg)) + msg


def pb_header(command: int, originator: str | None = None) -> bytes:
    m = pb_field_varint(1, command)
    if originator is not None:
        m += pb_field_len(2, originator.encode())
    return pb_delimited(m)


def pb_metadata(name: str, value: str) -> bytes:
    return pb_field_len(1, name.encode()) + pb_field_len(2, value.encode())


def pb_location_record(token: str, lat: float, lon: float, elev=None, date=None, metadata=None,
                       update_state=None, originator=None) -> bytes:
    m = pb_field_len(1, token.encode()) + pb_field_f64(2, lat) + pb_field_f64(3, lon)
    if elev is not None:
        m += pb_field_f64(4, elev)
    if date is not None:
        m += pb_field_fixed64(5, date)
    for k, v in (metadata or {}).items():
        m += pb_field_len(6, pb_metadata(k, v))
    if update_state is not None:
        m += pb_field_varint(7, int(update_state))
    return pb_header(3, originator) + pb_delimited(m)


def pb_alert_record(token: str, atype: str, message: str, date=None, metadata=None, update_state=None,
                    originator=None) -> bytes:
    m = pb_field_len(1, token.encode()) + pb_field_len(2, atype.encode()) + pb_field_len(3, message.encode())
    if date is not None:
        m += pb_field_fixed64(4, date)
    for k, v in (metadata or {}).items():
        m += pb_field_len(5, pb_metadata(k, v))
    if update_state is not None:
        m += pb_field_varint(6, int(update_state))
    return pb_header(4, originator) + pb_delimited(m)


def pb_measurements_record(token: str, mxs, date=None, metadata=None, update_state=None,
                           originator=None) -> bytes:
    m = pb_field_len(1, token.encode())
    for name, value in mxs:
        m += pb_field_len(2, pb_field_len(1, name.encode()) + pb_field_f64(2, value))
    if date is not None:
        m += pb_field_fixed64(3, date)
    for k, v in (metadata or {}).items():
        m += pb_field_len(4, pb_metadata(k, v))
    if update_state is not None:
        m += pb_field_varint(5, int(update_state))
    return pb_header(5, originator) + pb_delimited(m)
