"""Oracle: Kafka murmur2 known answers (kafka-clients UtilsTest.testMurmur2 vectors) and the
device-state fold against a plain Python dict fold.  CPU only."""
import numpy as np

import oracle
from sitewhere_b200 import synth as S

KAFKA_VECTORS = {b"21": -973932308, b"foobar": -790332482, b"a-little-bit-long-string": -985981536,
                 b"a-little-bit-longer-string": -1486304829,
                 b"lkjh234lh9fiuh90y23oiuhsafujhadof229phr9h19h89h8": -58897971, b"abc": 479470107}


def test_murmur2_kafka_vectors():
    for k, v in KAFKA_VECTORS.items():
        assert oracle.murmur2(k) == v & 0xFFFFFFFF
        assert oracle.partition(k, 7) == (v & 0x7FFFFFFF) % 7
        assert oracle.hash64(k) & 0xFFFFFFFF == v & 0xFFFFFFFF
    assert oracle.murmur2(b"") == oracle.murmur2(b"")  # length 0 handled


def python_fold(data, rec_off, cols, n, seq_base, id_hi, id_lo, state):
    raw = data.tobytes()
    for r in range(n):
        m = int(cols.meta[r])
        kind, fl = m & 0xFF, (m >> 8) & 0xFFFF
        if kind > 2 or fl & 32:
            continue
        rec = raw[int(rec_off[m >> 32]): int(rec_off[(m >> 32) + 1])]
        ts, ns = int(cols.token_span[r]), int(cols.name_span[r])
        tok = rec[ts & 0xFFFFFFFF:][: ts >> 32]
        name = b"" if kind == 1 else rec[ns & 0xFFFFFFFF:][: ns >> 32]
        kb = kind | (0x80 if (kind != 1 and not fl & 1024) else 0)
        key = len(tok).to_bytes(2, "little") + bytes([kb]) + tok + name
        state[key] = (seq_base + r, int(id_hi[r]), int(id_lo[r]), int(cols.event_date[r]), float(cols.v0[r]))


def test_state_fold_matches_python_dict():
    st = oracle.State()
    ref = {}
    seq = 0
    for b in range(3):
        data, rec_off, _ = S.pb_mixed(4000, 50, seed=100 + b)
        cols, status, first_row, total = oracle.decode_pb(data, rec_off)
        id_hi = np.full(total, b, np.uint64)
        id_lo = np.arange(total, dtype=np.uint64)
        st.merge(data, rec_off, cols, seq, id_hi, id_lo)
        python_fold(data, rec_off, cols, total, seq, id_hi, id_lo, ref)
        seq += total
    snap = st.snapshot()
    keys = sorted(ref)
    assert len(keys) == snap["seq"].shape[0]
    kb = snap["key_bytes"].tobytes()
    for i, k in enumerate(keys):
        assert kb[int(snap["key_off"][i]): int(snap["key_off"][i + 1])] == k
        s, hi, lo, d, v = ref[k]
        assert (int(snap["seq"][i]), int(snap["id_hi"][i]), int(snap["id_lo"][i]), int(snap["event_date"][i])) == (s, hi, lo, d)
        assert snap["value"][i] == v
    # 50 devices x (1 location + 16 names + 8 alert types) is the key-space bound
    assert len(keys) <= 50 * 25


def test_alt_id_set_known_answers():
    """oracle.AltIdSet (the restatement of AlternateIdDeduplicator as a set probe): first arrival is new, every later
    one is seen, requests without an alternate id are never duplicates, escaped ids are left to the store"""
    import oracle
    from sitewhere_b200 import synth as S
    mk = lambda alt: ('{"deviceToken":"t","type":"DeviceLocation","request":{"latitude":1,"alternateId":%s}}' % alt).encode()
    recs = [mk('"a"'), mk('"b"'), mk('"a"'), S.json_location("t", "1", "2", "3", 5), mk('"es\\u0063"'), mk('"b"'), mk('""'), mk('""')]
    data, off = S.pack_records(recs)
    cols, status, first, total = oracle.decode_json(data, off)
    assert total == len(recs)
    st = oracle.AltIdSet()
    assert st.probe(data, off, cols, total).tolist() == [0, 0, 1, 0, 2, 1, 0, 1]
    assert st.probe(data, off, cols, total).tolist() == [1, 1, 1, 0, 2, 1, 1, 1]
