"""The C-ABI library loads and exports every entry point include/swgpu.h declares (no compute)."""
import os
import re

from sitewhere_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "swgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(swgpu_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from sitewhere_b200 import build
    build.build()
    lib = L.load()
    syms = declared_symbols()
    assert len(syms) >= 27
    for s in syms:
        assert hasattr(lib, s), f"libswgpu.so does not export {s}"
        assert s in L.SIGNATURES, f"python binding lacks a signature for {s}"
    assert set(L.SIGNATURES) == set(syms)
    assert lib.swgpu_abi_version() == 2


def test_host_partition_helper_matches_kafka_vectors():
    import sitewhere_b200 as sw
    vec = {b"21": -973932308, b"foobar": -790332482, b"a-little-bit-long-string": -985981536,
           b"a-little-bit-longer-string": -1486304829, b"lkjh234lh9fiuh90y23oiuhsafujhadof229phr9h19h89h8": -58897971,
           b"abc": 479470107}
    for k, v in vec.items():
        assert sw.murmur2(k) == v & 0xFFFFFFFF
        assert sw.partition(k, 8) == ((v & 0x7FFFFFFF) % 8)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sitewhere_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "sw_oracle" not in src, f
