"""bench.py --impl reference on a tiny sample (CPU only): the JSON line carries the keys the driver reads, for every
workload; and the generators of all workloads produce what the oracle decodes cleanly."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("workload", ["c1", "c2", "c3", "c5", "n1"])
def test_reference_arm_line(workload):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", workload,
                          "--cpu-events", "3000", "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=600,
                         cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["higher_is_better"] is True and line["unit"] == "events/s"
    assert line["value"] > 0 and line["n_gpus"] == 1 and line["data"] == "synthetic"
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"]


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--cpu-events", "1000",
                          "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_enriched_generator_is_what_the_oracle_decodes():
    import oracle
    from sitewhere_b200 import synth as S
    lon = np.linspace(-85.0, -84.0, 500)
    lat = np.linspace(33.0, 35.0, 500)
    d, off, truth = S.enriched_locations(500, 40, lon, lat, seed=3)
    cols, ids, status, first, total = oracle.decode_enriched(d, off)
    assert total == 500 and not status.any()
    assert np.array_equal(cols.v0[:500], truth["lat"]) and np.array_equal(cols.v1[:500], truth["lon"])
    assert np.array_equal(cols.event_date[:500], truth["date"]) and np.array_equal(ids.event_hi[:500], truth["event_id"])
