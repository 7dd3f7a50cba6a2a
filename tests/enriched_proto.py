"""google.protobuf message classes for the enriched-event payload, from a hand-built descriptor of
sitewhere-grpc-event-management/src/main/proto/device-event-model.proto:139-215,288-296,321-324 and
sitewhere-grpc-model/src/main/proto/sitewhere-common.proto:12-33,146-151 (only the messages on the path).
An independent implementation of the wire format to pin the oracle with, and the generator of test payloads."""
import functools
import random
import struct


@functools.lru_cache(maxsize=1)
def classes():
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name = "enriched.proto"
    fdp.package = "com.sitewhere.grpc.model"
    fdp.syntax = "proto3"
    F = descriptor_pb2.FieldDescriptorProto
    P = ".com.sitewhere.grpc.model."

    def enum(name, values):
        e = fdp.enum_type.add()
        e.name = name
        for i, v in enumerate(values):
            x = e.value.add()
            x.name, x.number = v, i

    def msg(name, fields, oneof=None, maps=()):
        m = fdp.message_type.add()
        m.name = name
        if oneof:
            m.oneof_decl.add().name = oneof
        for (fname, num, ftype, tname) in fields:
            f = m.field.add()
            f.name, f.number, f.type = fname, num, ftype
            f.label = F.LABEL_REPEATED if fname in maps else F.LABEL_OPTIONAL
            if tname:
                f.type_name = P + tname
            if oneof:
                f.oneof_index = 0
        for mp in maps:        # map<string,string>: a nested *Entry message with map_entry = true
            ent = m.nested_type.add()
            ent.name = "".join(w.capitalize() if i else w[0].upper() + w[1:] for i, w in enumerate([mp])) + "Entry"
            ent.options.map_entry = True
            for (fname, num) in (("key", 1), ("value", 2)):
                f = ent.field.add()
                f.name, f.number, f.type, f.label = fname, num, F.TYPE_STRING, F.LABEL_OPTIONAL
            for f in m.field:
                if f.name == mp:
                    f.type_name = P + name + "." + ent.name
        return m

    enum("GDeviceEventType", ["EVENT_TYPE_MEASUREMENT", "EVENT_TYPE_LOCATION", "EVENT_TYPE_ALERT",
                              "EVENT_TYPE_COMMAND_INVOCATION", "EVENT_TYPE_COMMAND_RESPONSE", "EVENT_TYPE_STATE_CHANGE"])
    enum("GAlertSource", ["ALERT_SOURCE_DEVICE", "ALERT_SOURCE_SYSTEM"])
    enum("GAlertLevel", ["ALERT_LEVEL_INFO", "ALERT_LEVEL_WARNING", "ALERT_LEVEL_ERROR", "ALERT_LEVEL_CRITICAL"])
    enum("GDeviceAssignmentStatus", ["ASSN_STATUS_UNSPECIFIED", "ASSN_STATUS_ACTIVE", "ASSN_STATUS_MISSING",
                                     "ASSN_STATUS_RELEASED"])
    msg("GOptionalDouble", [("value", 1, F.TYPE_DOUBLE, "")])
    msg("GOptionalString", [("value", 1, F.TYPE_STRING, "")])
    msg("GUUID", [("msb", 1, F.TYPE_FIXED64, ""), ("lsb", 2, F.TYPE_FIXED64, "")])
    msg("GDeviceEvent", [("id", 1, F.TYPE_MESSAGE, "GUUID"), ("alternateId", 2, F.TYPE_MESSAGE, "GOptionalString"),
                         ("eventType", 3, F.TYPE_ENUM, "GDeviceEventType"), ("deviceId", 4, F.TYPE_MESSAGE, "GUUID"),
                         ("deviceAssignmentId", 5, F.TYPE_MESSAGE, "GUUID"), ("customerId", 6, F.TYPE_MESSAGE, "GUUID"),
                         ("areaId", 7, F.TYPE_MESSAGE, "GUUID"), ("assetId", 8, F.TYPE_MESSAGE, "GUUID"),
                         ("eventDate", 9, F.TYPE_UINT64, ""), ("receivedDate", 10, F.TYPE_UINT64, ""),
                         ("metadata", 11, F.TYPE_MESSAGE, "")], maps=("metadata",))
    msg("GDeviceMeasurement", [("name", 1, F.TYPE_STRING, ""), ("value", 2, F.TYPE_DOUBLE, ""),
                               ("event", 3, F.TYPE_MESSAGE, "GDeviceEvent")])
    msg("GDeviceAlert", [("source", 1, F.TYPE_ENUM, "GAlertSource"), ("level", 2, F.TYPE_ENUM, "GAlertLevel"),
                         ("type", 3, F.TYPE_STRING, ""), ("alertMessage", 4, F.TYPE_STRING, ""),
                         ("event", 5, F.TYPE_MESSAGE, "GDeviceEvent")])
    msg("GDeviceLocation", [("latitude", 1, F.TYPE_MESSAGE, "GOptionalDouble"), ("longitude", 2, F.TYPE_MESSAGE, "GOptionalDouble"),
                            ("elevation", 3, F.TYPE_MESSAGE, "GOptionalDouble"), ("event", 4, F.TYPE_MESSAGE, "GDeviceEvent")])
    msg("GDeviceStateChange", [("attribute", 1, F.TYPE_STRING, ""), ("type", 2, F.TYPE_STRING, ""),
                               ("event", 5, F.TYPE_MESSAGE, "GDeviceEvent")])
    msg("GOtherEvent", [])      # stands for the event types off the path: checked at the wire level only
    msg("GAnyDeviceEvent", [("measurement", 1, F.TYPE_MESSAGE, "GDeviceMeasurement"), ("alert", 2, F.TYPE_MESSAGE, "GDeviceAlert"),
                            ("location", 3, F.TYPE_MESSAGE, "GDeviceLocation"), ("commandInvocation", 4, F.TYPE_MESSAGE, "GOtherEvent"),
                            ("commandResponse", 5, F.TYPE_MESSAGE, "GOtherEvent"),
                            ("stateChange", 6, F.TYPE_MESSAGE, "GOtherEvent")], oneof="event")
    msg("GDeviceEventContext", [("deviceId", 1, F.TYPE_MESSAGE, "GUUID"), ("deviceTypeId", 2, F.TYPE_MESSAGE, "GUUID"),
                                ("parentDeviceId", 3, F.TYPE_MESSAGE, "GUUID"), ("deviceStatus", 4, F.TYPE_MESSAGE, "GOptionalString"),
                                ("deviceMetadata", 5, F.TYPE_MESSAGE, ""), ("assignmentStatus", 6, F.TYPE_ENUM, "GDeviceAssignmentStatus"),
                                ("assignmentMetadata", 7, F.TYPE_MESSAGE, "")], maps=("deviceMetadata", "assignmentMetadata"))
    msg("GEnrichedEventPayload", [("context", 1, F.TYPE_MESSAGE, "GDeviceEventContext"), ("event", 2, F.TYPE_MESSAGE, "GAnyDeviceEvent")])
    # the outbound side (device-event-model.proto:34-42,45-49,66-72,75-80,111-120,307-312; sitewhere-common.proto GOptionalBoolean)
    msg("GOptionalBoolean", [("value", 1, F.TYPE_BOOL, "")])
    msg("GDeviceEventCreateRequest", [("alternateId", 1, F.TYPE_MESSAGE, "GOptionalString"),
                                      ("customerToken", 2, F.TYPE_MESSAGE, "GOptionalString"),
                                      ("areaToken", 3, F.TYPE_MESSAGE, "GOptionalString"),
                                      ("assetToken", 4, F.TYPE_MESSAGE, "GOptionalString"), ("eventDate", 5, F.TYPE_UINT64, ""),
                                      ("updateState", 6, F.TYPE_MESSAGE, "GOptionalBoolean"), ("metadata", 7, F.TYPE_MESSAGE, "")],
        maps=("metadata",))
    msg("GDeviceMeasurementCreateRequest", [("name", 1, F.TYPE_STRING, ""), ("value", 2, F.TYPE_DOUBLE, ""),
                                            ("event", 3, F.TYPE_MESSAGE, "GDeviceEventCreateRequest")])
    msg("GDeviceAlertCreateRequest", [("source", 1, F.TYPE_ENUM, "GAlertSource"), ("level", 2, F.TYPE_ENUM, "GAlertLevel"),
                                      ("type", 3, F.TYPE_STRING, ""), ("alertMessage", 4, F.TYPE_STRING, ""),
                                      ("event", 5, F.TYPE_MESSAGE, "GDeviceEventCreateRequest")])
    msg("GDeviceLocationCreateRequest", [("latitude", 1, F.TYPE_MESSAGE, "GOptionalDouble"),
                                         ("longitude", 2, F.TYPE_MESSAGE, "GOptionalDouble"),
                                         ("elevation", 3, F.TYPE_MESSAGE, "GOptionalDouble"),
                                         ("event", 4, F.TYPE_MESSAGE, "GDeviceEventCreateRequest")])
    msg("GAnyDeviceEventCreateRequest", [("measurement", 1, F.TYPE_MESSAGE, "GDeviceMeasurementCreateRequest"),
                                         ("alert", 2, F.TYPE_MESSAGE, "GDeviceAlertCreateRequest"),
                                         ("location", 3, F.TYPE_MESSAGE, "GDeviceLocationCreateRequest")], oneof="event")
    msg("GInboundEventPayload", [("sourceId", 1, F.TYPE_STRING, ""), ("deviceToken", 2, F.TYPE_STRING, ""),
                                 ("originator", 3, F.TYPE_MESSAGE, "GOptionalString"),
                                 ("event", 4, F.TYPE_MESSAGE, "GAnyDeviceEventCreateRequest")])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("com.sitewhere.grpc.model." + n))
    return {n: get(n) for n in ("GEnrichedEventPayload", "GAnyDeviceEvent", "GDeviceEvent", "GDeviceMeasurement", "GDeviceAlert",
                                "GDeviceLocation", "GDeviceEventContext", "GUUID", "GOptionalDouble", "GOptionalString",
                                "GInboundEventPayload", "GOptionalBoolean")}


NAMES = ["fuel.level", "engine.temp", "oil.psi", "rpm.avg", "speed.kmh", "battery.v", "", "température"]
ALERTS = ["engine.overheat", "fuel.low", "geofence.exit", "x"]


def _uuid(m, rnd, base=0):
    m.msb = rnd.getrandbits(64) if base == 0 else base
    m.lsb = rnd.getrandbits(64)


def random_payload(rnd: random.Random, n_assign=50, kinds=(1, 2, 3), sparse=0.15, neg_zero=True):
    """one well-formed GEnrichedEventPayload (python message object)"""
    C = classes()
    p = C["GEnrichedEventPayload"]()
    _uuid(p.context.deviceId, rnd)
    _uuid(p.context.deviceTypeId, rnd)
    if rnd.random() < 0.3:
        p.context.deviceStatus.value = "ok"
    if rnd.random() < 0.3:
        p.context.deviceMetadata["fw"] = "1.%d" % rnd.randrange(9)
    p.context.assignmentStatus = rnd.randrange(4)
    kind = rnd.choice(kinds)
    if kind == 1:
        ev = p.event.measurement
        if rnd.random() > sparse:
            ev.name = rnd.choice(NAMES)
        if rnd.random() > sparse:
            ev.value = rnd.choice([0.0, -0.0 if neg_zero else 2.5, 1.5, rnd.uniform(-1e3, 1e3), float("inf")])
    elif kind == 2:
        ev = p.event.alert
        ev.source = rnd.randrange(2)
        ev.level = rnd.randrange(4)
        if rnd.random() > sparse:
            ev.type = rnd.choice(ALERTS)
        if rnd.random() > sparse:
            ev.alertMessage = "alert " + str(rnd.randrange(1000))
    elif kind == 3:
        ev = p.event.location
        if rnd.random() > sparse:
            ev.latitude.value = rnd.choice([0.0, rnd.uniform(34, 35)])
        if rnd.random() > sparse:
            ev.longitude.value = rnd.uniform(-85, -84)
        if rnd.random() > 0.5:
            ev.elevation.value = rnd.uniform(0, 500)
    else:
        p.event.stateChange.SetInParent()
        return p
    e = ev.event
    _uuid(e.id, rnd)
    a = rnd.randrange(n_assign)
    e.deviceAssignmentId.msb = 0x1000 + a
    e.deviceAssignmentId.lsb = 0xABCDEF0000000000 + a * 7
    _uuid(e.deviceId, rnd)
    e.eventType = {1: 0, 2: 2, 3: 1}.get(kind, 5) if rnd.random() > 0.05 else rnd.randrange(6)
    if rnd.random() > sparse:
        e.eventDate = 1_527_854_400_000 + rnd.randrange(10 ** 9)
    e.receivedDate = 1_527_854_400_000
    if rnd.random() < 0.2:
        e.alternateId.value = "alt-%d" % rnd.randrange(100) if rnd.random() < 0.8 else ""
    if rnd.random() < 0.2:
        e.metadata["k%d" % rnd.randrange(3)] = "v"
    return p


def expected_row(rec: bytes):
    """(status, dict) for one record as the reference's parse + conversion gives it, via google.protobuf"""
    C = classes()
    p = C["GEnrichedEventPayload"]()
    try:
        p.ParseFromString(rec)
    except Exception:
        return 1, None
    which = p.event.WhichOneof("event")
    if which in ("stateChange", "commandInvocation", "commandResponse"):
        return 3, None
    if p.context.assignmentStatus > 3 or which is None:
        return 1, None
    ev = getattr(p.event, which)
    e = ev.event
    if e.eventType > 5 or (which == "alert" and (ev.source > 1 or ev.level > 3)):
        return 1, None
    d = dict(kind={"measurement": 0, "location": 1, "alert": 2}[which], etype=e.eventType,
             date=e.eventDate, event_id=(e.id.msb, e.id.lsb), assignment=(e.deviceAssignmentId.msb, e.deviceAssignmentId.lsb),
             device=(e.deviceId.msb, e.deviceId.lsb), has_alt=e.HasField("alternateId"), alt=e.alternateId.value,
             n_meta=len(e.metadata))
    bits = lambda x: struct.unpack("<Q", struct.pack("<d", x))[0]
    if which == "measurement":
        d.update(name=ev.name, v0=bits(ev.value))
    elif which == "alert":
        d.update(name=ev.type, message=ev.alertMessage, level=ev.level, source=ev.source)
    else:
        d.update(has_lat=ev.HasField("latitude"), has_lon=ev.HasField("longitude"), has_elev=ev.HasField("elevation"),
                 v0=bits(ev.latitude.value), v1=bits(ev.longitude.value), v2=bits(ev.elevation.value))
    return 0, d
