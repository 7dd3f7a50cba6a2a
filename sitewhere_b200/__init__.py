"""sitewhere_b200 — B200 (sm_100a) implementation of SiteWhere's inbound event hot path:
payload decode -> geofence zone test -> device-state merge, behind the C ABI in include/swgpu.h.
There is no CPU path: importing is cheap, but every operation needs libswgpu.so and a B200."""
from . import _lib  # noqa: F401
from .api import (Context, DeviceCols, HostCols, DecodeResult, EventDecodeException, DecoderRuntimeException,
                  PassThrough, ProtobufDeviceEventDecoder, JsonDeviceRequestDecoder, ZoneTest,
                  ZoneTestRuleProcessor, DeviceStateProcessingLogic, murmur2, partition)  # noqa: F401

__all__ = ["Context", "DeviceCols", "HostCols", "DecodeResult", "EventDecodeException", "DecoderRuntimeException",
           "PassThrough", "ProtobufDeviceEventDecoder", "JsonDeviceRequestDecoder", "ZoneTest",
           "ZoneTestRuleProcessor", "DeviceStateProcessingLogic", "murmur2", "partition"]
