#!/usr/bin/env python
"""bench.py — events/sec of the inbound hot path on B200, one line of JSON per run.

Workloads (BASELINE.json `configs`; --workload, default c2 = the configuration the metric is quoted on):
  c1  1 M JSON DeviceMeasurement requests, JsonDeviceRequestDecoder + zone test vs 16 rectangular zones
  c2  10 M protobuf DeviceLocation events per GPU, ProtobufDeviceEventDecoder + zone test vs 1,024 concave
      64-vertex zones
  c3  100 M mixed measurement/location/alert protobuf events over 1 M devices, decode + device-state merge,
      in batches of 10 M records with the sequence base advancing
  c5  a stream of 64 KB polls (about 1,040 location events each), 10 M devices, 65,536 zones: sustained
      events/s and per-poll latency through the small-batch path of swgpu_decode_zone_test
A step is one pass of the path over the workload's synthetic input.

  value        events/s, inputs resident in HBM, CUDA events on the launch stream, max over ranks
  e2e          events/s through the host-buffer C-ABI call on pinned host memory (H2D + D2H inside the timed region)
  roofline     the dominant kernel's algorithmic GB/s (SURVEY.md §8(d) byte counts) vs the measured HBM peak
  cpu_baseline the CPU oracle (C restatement of the JVM path) on a bounded sample of the same input, all host cores

  python bench.py [--workload c1|c2|c3|c5] [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
Multi-GPU: torchrun launches one rank per GPU; events are sharded by Kafka's partitioner on the device token,
the zone table is built on rank 0 and replicated with one NCCL broadcast; there is no per-batch collective
(weak scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

UNIT = "events/s"
N_DEVICES = 1_000_000

WORKLOADS = {
    "c1": dict(metric="events/sec decode+zone-test",
               name="C1: 1M JSON DeviceMeasurement requests, JsonDeviceRequestDecoder + zone test vs 16 rectangular zones",
               events=1_000_000),
    "c2": dict(metric="events/sec decode+zone-test",
               name="C2: 10M protobuf DeviceLocation decode + zone test vs 1,024 concave 64-vertex zones",
               events=10_000_000),
    "c3": dict(metric="events/sec decode+merge",
               name="C3: 100M mixed measurement/location/alert protobuf events, 1M device tokens, decode + device-state merge",
               events=100_000_000),
    "n1": dict(metric="events/sec decode+zone-test",
               name="N1 (SURVEY.md section 8f): 10M GEnrichedEventPayload location records off the inbound-enriched-events topic, "
                    "enriched decode + zone test vs 1,024 concave 64-vertex zones; decode + device-state merge beside it",
               events=10_000_000),
    "c5": dict(metric="events/sec decode+zone-test (64 KB micro-batch stream)",
               name="C5: 64 KB micro-batch stream of protobuf DeviceLocation events, 10M devices, 65,536 concave 64-vertex zones",
               events=2_000_000),
}
C3_BATCH = 10_000_000
C5_POLL_BYTES = 65_536
CPU_SAMPLE = {"c1": 300_000, "c2": 2_000_000, "c3": 4_000_000, "c5": 24_000, "n1": 2_000_000}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def fp64_peak():
    """Measured DFMA peak: tools/exp/dfma run now if it is there (our own microbenchmark), else the committed run."""
    exe = os.path.join(ROOT, "tools", "exp", "dfma")
    try:
        if os.path.exists(exe):
            out = subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout.strip().splitlines()[-1]
            d = json.loads(out)
            if d.get("dfma_tflops"):
                return float(d["dfma_tflops"]), "measured now (tools/exp/dfma)"
    except Exception:
        pass
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_dfma.json")))
        return float(d["dfma_tflops"]), "measured earlier (profiles/r02_dfma.json)"
    except Exception:
        return None, "unmeasured"


def profile_traffic(kernel: str, units: int):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture (profiles/traffic.json records the
    commit it was taken at); None when there is no capture of this kernel at this size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        for e in t["entries"]:
            if e["kernel"] == kernel and int(e["units"]) == int(units):
                return {"dram_bytes": e["dram_bytes"], "captured_at_commit": t.get("commit"), "source": e.get("source")}
    except Exception:
        pass
    return None


def random_access_ceiling(ms_fused: float, ms_unfused_merge: float, rows: int):
    """The merge's table accesses are random sectors, not a stream: tools/exp/randacc.cu measures that access pattern
    alone (random 64-byte read + atomicMax + one 256-bit store per row, the merge's own instructions; profiles/r02g_randacc.jsonl, largest table).
    Returns the measured ceiling and how close the kernels are to it; None without the file."""
    try:
        best = None
        for l in open(os.path.join(ROOT, "profiles", "r02d_randacc.jsonl")):
            e = json.loads(l)
            if e.get("variant") == "read64+atomicMax+store32 (one 256-bit store)" and (best is None or e["table_MB"] > best["table_MB"]):
                best = e
        ms = best["ms_per_10M_rows"] * rows / 1e7
        return {"source": "profiles/r02g_randacc.jsonl (tools/exp/randacc.cu, %d MB table)" % best["table_MB"],
                "ms_per_batch": ms, "frac_unfused_merge": ms / ms_unfused_merge, "frac_fused_decode_merge": ms / ms_fused,
                "note": "fractions = ceiling time / kernel time; the fused kernel also decodes the payload in that time"}
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.samples = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        for s in self.samples:
            try:
                sm.append(float(s[1]))
                mx.append(float(s[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ inputs
def zones_for(workload: str):
    from sitewhere_b200 import synth as S
    if workload == "c1":
        return S.rect_zones(4)
    if workload == "c5":
        return S.star_zones(256, nvert=64, seed=S.SEED_BASE + 4)
    return S.star_zones(32, nvert=64, seed=S.SEED_BASE + 1)


def gen_locations(n, zones, n_devices, rank, world, seed):
    from sitewhere_b200 import synth as S
    lon, lat = S.locations_for_zones(n, *zones[:3], seed=seed + 1000 * rank)
    devices = S.devices_of_shard(n_devices, rank, world) if world > 1 else None
    data, rec_off, _ = S.pb_locations(n, n_devices, lon, lat, seed=seed + 1000 * rank, devices=devices)
    return data, rec_off


def gen_n1(n, zones, rank):
    from sitewhere_b200 import synth as S
    lon, lat = S.locations_for_zones(n, *zones[:3], seed=S.SEED_BASE + 9 + 1000 * rank)
    data, rec_off, _ = S.enriched_locations(n, N_DEVICES, lon, lat, seed=S.SEED_BASE + 9 + 1000 * rank)
    return data, rec_off


def gen_c1(n, rank):
    """n JSON measurement requests; the generator's Python loop is slow, so 250 K distinct payloads are tiled"""
    from sitewhere_b200 import synth as S
    base = min(n, 250_000)
    data, rec_off, _ = S.json_measurements(base, 100_000, seed=S.SEED_BASE + 77 * rank)
    reps = (n + base - 1) // base
    if reps > 1:
        lens = np.tile(np.diff(rec_off.astype(np.int64)), reps)[:n]
        data = np.tile(data, reps)[: int(lens.sum())]
        rec_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    return data, rec_off


def gen_c3_batch(b, n, rank):
    from sitewhere_b200 import synth as S
    data, rec_off, _ = S.pb_mixed(n, N_DEVICES, seed=S.SEED_BASE + 2 + 101 * b + 7919 * rank)
    return data, rec_off


def build_inputs(workload, n, rank, world):
    from sitewhere_b200 import synth as S
    zones = zones_for(workload) if workload != "c3" else None
    if workload == "c1":
        return zones, gen_c1(n, rank)
    if workload == "c2":
        return zones, gen_locations(n, zones, N_DEVICES, rank, world, S.SEED_BASE + 1)
    if workload == "n1":
        return zones, gen_n1(n, zones, rank)
    if workload == "c5":
        return zones, gen_locations(n, zones, 10_000_000, rank, world, S.SEED_BASE + 4)
    nb = max(1, (n + C3_BATCH - 1) // C3_BATCH)
    per = n // nb
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(min(nb, max(1, (os.cpu_count() or 4) // 2))) as ex:
        batches = list(ex.map(lambda b: gen_c3_batch(b, per, rank), range(nb)))
    return None, batches


# ------------------------------------------------------------------------------------------------ CPU legs (oracle)
def _threads_map(fn, n_items, threads):
    from concurrent.futures import ThreadPoolExecutor
    bounds = np.linspace(0, n_items, threads + 1).astype(np.int64)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda k: fn(int(bounds[k]), int(bounds[k + 1])), range(threads)))
    return time.perf_counter() - t0


def cpu_decode_zone(fmt, zones, data, rec_off, threads):
    """The oracle's decode + index-free zone loop, records split over `threads` host threads (ctypes drops the GIL)."""
    import oracle
    oracle.lib()
    def dec(chunk, off, max_rows):
        if fmt == "enriched":
            cols, ids, status, first_row, total = oracle.decode_enriched(chunk, off, max_rows=max_rows)
            return cols, status, first_row, total
        return (oracle.decode_json if fmt == "json" else oracle.decode_pb)(chunk, off, max_rows=max_rows)

    def work(a, b):
        if a == b:
            return
        off = rec_off[a: b + 1] - rec_off[a]
        chunk = data[int(rec_off[a]): int(rec_off[b])]
        cols, status, first_row, total = dec(chunk, off, max_rows=b - a)
        oracle.zone_test(*zones, cols.v1[:total], cols.v0[:total], meta=cols.meta[:total], max_pairs=max(total, 1))
    return _threads_map(work, rec_off.shape[0] - 1, threads)


def cpu_decode_merge(data, rec_off, threads):
    """The oracle's decode + sequential fold; each thread folds its slice into its own state (as one consumer per
    Kafka partition would)."""
    import oracle
    oracle.lib()

    def work(a, b):
        if a == b:
            return
        off = rec_off[a: b + 1] - rec_off[a]
        chunk = data[int(rec_off[a]): int(rec_off[b])]
        cols, status, first_row, total = oracle.decode_pb(chunk, off, max_rows=b - a)
        ids = np.arange(total, dtype=np.uint64)
        st = oracle.State()
        st.merge(chunk, off, cols, 0, ids, ids)
    return _threads_map(work, rec_off.shape[0] - 1, threads)


def cpu_leg(workload, zones, rank_inputs, cores, sample):
    """Times the oracle on the first `sample` events of this rank's input.  Returns (events, seconds, description)."""
    if workload == "c3":
        data, rec_off = rank_inputs[0]
        sample = min(sample, rec_off.shape[0] - 1)
        t = cpu_decode_merge(data[: int(rec_off[sample])], rec_off[: sample + 1], cores)
        return sample, t, f"first {sample} events of batch 0: decode + sequential fold, one state per thread"
    data, rec_off = rank_inputs
    sample = min(sample, rec_off.shape[0] - 1)
    fmt = "json" if workload == "c1" else "enriched" if workload == "n1" else "pb"
    t = cpu_decode_zone(fmt, zones, data[: int(rec_off[sample])], rec_off[: sample + 1], cores)
    return sample, t, f"first {sample} events of the same input: decode + index-free zone loop over {len(zones[3])} zones"


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The JVM path cannot run here (no JDK /
    jars), so this times the CPU oracle (a C restatement, kind 'port') on all host cores, on a bounded sample of
    the same workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    w = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    sample = args.cpu_events or CPU_SAMPLE[args.workload]
    zones, inputs = build_inputs(args.workload, sample, 0, 1)
    for _ in range(max(1, args.warmup // 3)):
        cpu_leg(args.workload, zones, inputs, cores, max(sample // 10, 1000))
    t, ev, desc = 0.0, 0, ""
    for _ in range(args.steps):
        e, dt, desc = cpu_leg(args.workload, zones, inputs, cores, sample)
        t += dt
        ev += e
    v = ev / t
    line = {"metric": w["metric"], "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64", "data": "synthetic",
            "impl": "reference",
            "config": {"workload": w["name"], "sample_events_per_step": ev // args.steps},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": desc + " (C oracle; JVM path unavailable: no JDK on this box)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ link probe
def pcie_probe(torch, dist, dev, world, nbytes=256 << 20, reps=4):
    """Host<->device copy rates of this box (pinned memory), all ranks copying AT THE SAME TIME behind a barrier:
    each direction alone, then both at once.  Returns this rank's rates and, for N > 1, the sum over ranks —
    the ceiling of the end-to-end number."""
    hin = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    hout = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    din = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dout = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(up, down):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        if up:
            ev[0].record(s1)
        if down:
            ev[2].record(s2)
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s1):
                    din.copy_(hin, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    hout.copy_(dout, non_blocking=True)
        if up:
            ev[1].record(s1)
        if down:
            ev[3].record(s2)
        torch.cuda.synchronize()
        gb = nbytes * reps / 1e6
        return (gb / ev[0].elapsed_time(ev[1]) if up else 0.0, gb / ev[2].elapsed_time(ev[3]) if down else 0.0)

    run(True, True)
    h, _ = run(True, False)
    _, d = run(False, True)
    dh, dd = run(True, True)
    out = {"h2d_GBps": h, "d2h_GBps": d, "duplex_h2d_GBps": dh, "duplex_d2h_GBps": dd, "ranks_copying_at_once": world}
    if world > 1:
        tot = torch.tensor([h, d, dh, dd], dtype=torch.float64, device=dev)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        th, td, tdh, tdd = [float(x) for x in tot.tolist()]
        out["aggregate"] = {"h2d_GBps": th, "d2h_GBps": td, "duplex_h2d_GBps": tdh, "duplex_d2h_GBps": tdd}
    return out


# ------------------------------------------------------------------------------------------------ helpers
class Env:
    clocks = None


def stage_timer(torch, stream, nstages):
    """events around the stages of every timed step, recorded on the stream the kernels are launched on"""
    marks = []

    def begin():
        marks.append([torch.cuda.Event(enable_timing=True) for _ in range(nstages + 1)])
        marks[-1][0].record(stream)

    def mark(k):
        marks[-1][k].record(stream)

    def mean_ms():
        return [float(np.mean([m[k].elapsed_time(m[k + 1]) for m in marks])) for k in range(nstages)]
    return begin, mark, mean_ms


def to_dev(torch, dev, data, rec_off):
    d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
    d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
    d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
    return d_bytes, d_off


def timed_region(env, step, steps, warmup):
    """W warm-up steps, then K steps between barriers, CUDA events on the stream; returns this rank's ms and
    leaves the clocks sampled during the region in env.clocks (rank 0)"""
    torch = env.torch
    for _ in range(warmup):
        step(False)
    env.barrier()
    sampler = ClockSampler(env.local)
    if env.rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(env.stream)
    for _ in range(steps):
        step(True)
    e1.record(env.stream)
    env.barrier()
    ms = e0.elapsed_time(e1)
    # the timed region may last only milliseconds and nvidia-smi answers in ~100 ms: keep the same steps running
    # for about a second more so that the clock / throttle-reason samples are taken under the same load
    t_probe = time.perf_counter() + 1.2
    while time.perf_counter() < t_probe:
        for _ in range(4):
            step(False)
        torch.cuda.synchronize()
    if env.rank == 0:
        env.clocks = sampler.stop()
    return ms


def max_over_ranks(env, x):
    t = env.torch.tensor([x], dtype=env.torch.float64, device=env.dev)
    if env.world > 1:
        env.dist.all_reduce(t, op=env.dist.ReduceOp.MAX)
    return float(t.item())


def e2e_region(env, call, steps):
    for _ in range(2):
        call()
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    env.torch.cuda.synchronize()
    return max_over_ranks(env, time.perf_counter() - t0)


def zone_block(ctx, run_zone, ms_zone, n_loc):
    """the zone stage in its own units: work counters from one counted (untimed) pass / the timed stage"""
    ctx.set_option("zone.stats", 1)
    try:
        run_zone()
        st = ctx.zone_stats()
    finally:
        ctx.set_option("zone.stats", 0)
    peak, src = fp64_peak()
    flops = 12.0 * st["straddle_tests"] + 4.0 * st["edge_records"]
    sec = ms_zone * 1e-3
    return {"ms": ms_zone, "bound": "FP64 pipe / instruction issue (zone table L2-resident), not HBM",
            "locations_per_s": n_loc / sec, "cell_records_per_s": st["cell_records"] / sec,
            "edge_records_per_s": st["edge_records"] / sec, "straddle_tests_per_s": st["straddle_tests"] / sec,
            "edge_records_per_location": st["edge_records"] / max(n_loc, 1),
            "dfma_equiv_flops_per_step": flops, "dfma_equiv_tflops": flops / sec / 1e12,
            "fp64_peak_tflops": peak, "fp64_peak_source": src,
            "frac_fp64_peak": (flops / sec / 1e12 / peak) if peak else None,
            "flop_count": "12 per straddle test + 4 per edge record (SURVEY.md §8(d)), over the grid-indexed walk "
                          "(the index-free JTS loop would do T envelope tests per location first)",
            "note": "ms is the whole zone stage (zone_count_kernel + zone_scan_mc_kernel + zone_emit_kernel)"}


# ------------------------------------------------------------------------------------------------ c1 / c2
def run_decode_zone(args, env, workload):
    torch, sw, ctx, dev, stream = env.torch, env.sw, env.ctx, env.dev, env.stream
    from sitewhere_b200.api import pinned_array, pinned_cols
    fmt = "json" if workload == "c1" else "enriched" if workload == "n1" else "pb"
    zones, (data, rec_off) = build_inputs(workload, args.events, env.rank, env.world)
    n = rec_off.shape[0] - 1
    if env.rank == 0:
        ctx.set_zones(*zones)
    if env.world > 1:
        from sitewhere_b200.shard import broadcast_zone_table
        broadcast_zone_table(ctx, env.dist, dev, env.rank)
    total_bytes = int(rec_off[-1])
    d_bytes, d_off = to_dev(torch, dev, data, rec_off)
    cols = sw.DeviceCols(n, dev, optional=False)
    d_status = torch.empty(n, dtype=torch.uint8, device=dev)
    d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
    max_pairs = max(n, 1024)
    d_pairs = torch.empty(max_pairs, dtype=torch.int64, device=dev)
    d_npairs = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    decode = {"json": ctx.decode_json_dev, "enriched": ctx.decode_enriched_dev, "pb": ctx.decode_pb_dev}[fmt]
    begin, mark, mean_ms = stage_timer(torch, stream, 2)

    def run_zone():
        ctx.zone_test_dev(cols.v1, cols.v0, cols.meta, n, d_nrows, d_pairs, d_npairs)

    # the timed step is ONE device-buffer call, swgpu_decode_zone_test_dev (decode, then the zone test on the columns)
    def step(timed):
        ctx.decode_zone_test_dev(fmt, d_bytes, d_off, n, cols, d_status, d_first, d_nrows, d_pairs, d_npairs)

    # the same work as two calls (swgpu_decode_*_dev, then swgpu_zone_test_dev on the decoded columns), timed per
    # stage AFTER the timed region: what each stage costs on its own
    def step_two_calls(timed):
        if timed:
            begin()
        decode(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        if timed:
            mark(1)
        run_zone()
        if timed:
            mark(2)

    with torch.cuda.stream(stream):
        ms = timed_region(env, step, args.steps, args.warmup)
        n_rows, n_pairs = int(d_nrows.item()), int(d_npairs.item())
        assert n_rows == n, (n_rows, n)
        assert int(d_status.max().item()) == 0
        for k in range(args.warmup + args.steps):
            step_two_calls(k >= args.warmup)
        torch.cuda.synchronize()
        assert (int(d_nrows.item()), int(d_npairs.item())) == (n_rows, n_pairs)
        ms_dec, ms_zone = mean_ms()
        zb = zone_block(ctx, run_zone, ms_zone, n_rows) if env.rank == 0 else None
        # SURVEY.md §8 row N2 (not part of the timed step): the outbound GInboundEventPayload of every row, built on
        # the device from the columns just decoded
        enc = None
        mrg = None
        if env.rank == 0 and fmt == "enriched":
            # the device-state consumer's side of the same topic: decode + merge in one kernel, presence rows kept
            ctx.state_clear()
            ctx.state_reserve(4 * N_DEVICES, 1 << 20)
            seqn = [0]

            def run_mrg():
                ctx.set_option("state.presence_now_ms", 1_700_000_000_000 + seqn[0])
                ctx.decode_merge_dev("enriched", d_bytes, d_off, n, seqn[0], None, None, d_status, d_nrows)
                seqn[0] += n << 12
            for _ in range(3):
                run_mrg()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.steps):
                run_mrg()
            e1.record(stream)
            torch.cuda.synchronize()
            ms_m = e0.elapsed_time(e1) / args.steps
            keys, _kb = ctx.state_size()
            ctx.set_option("state.presence_now_ms", 0)
            mrg = {"ms": ms_m, "events_per_s": n / (ms_m * 1e-3), "state_keys": keys,
                   "what": "swgpu_decode_merge_dev(SWGPU_FMT_ENRICHED) with state.presence_now_ms set: one kernel parses the "
                           "payload, folds the location and touches the device's presence row"}
        if env.rank == 0 and fmt != "enriched":
            src = b"bench-source"
            d_eoff = torch.empty(n + 1, dtype=torch.int64, device=dev)
            d_est = torch.empty(n, dtype=torch.uint8, device=dev)
            d_etot = torch.zeros(1, dtype=torch.int64, device=dev)
            d_eout = torch.empty(int(total_bytes) + (48 + len(src)) * n, dtype=torch.uint8, device=dev)

            def run_enc():
                ctx.encode_inbound_dev(fmt, d_bytes, d_off, cols, n, d_nrows, src, 1_700_000_000_000, d_eout, d_eoff, d_est,
                                       d_etot)
            for _ in range(3):
                run_enc()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.steps):
                run_enc()
            e1.record(stream)
            torch.cuda.synchronize()
            ms_enc = e0.elapsed_time(e1) / args.steps
            out_b = int(d_etot.item())
            assert out_b <= d_eout.numel() and int(d_est.max().item()) == 0
            # bytes moved: the columns it reads (meta, date, three values, two spans), the strings through the spans
            # (token + name), sizes written and read back, offsets, status, and the records written
            moved = 56 * n + out_b + 4 * 2 * n + 8 * n + n + 32 * n
            enc = {"ms": ms_enc, "events_per_s": n / (ms_enc * 1e-3), "out_bytes": out_b,
                   "out_bytes_per_event": out_b / n, "approx_bytes_moved": moved,
                   "approx_frac_hbm": moved / (ms_enc * 1e-3) / 1e9 / measured_peaks()[0],
                   "launches": "encode_size_kernel, encode_scan_kernel, encode_emit_kernel"}
            del d_eout
    ms_all = max_over_ranks(env, ms)
    value = n * env.world * args.steps / (ms_all * 1e-3)

    # ---- end to end: ONE swgpu_decode_zone_test call per step on PINNED host buffers (what the Java binding's
    # staging segments are).  "full": every field of the decoded request comes back (meta, eventDate, the three
    # doubles, the token span; + name span for measurements), with status, first_row and the alert pairs.
    # "alerts_only": status + alert pairs (what the rule-processing shim needs).
    e2e = {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    if not args.no_e2e:
        e2e_cols = ["meta", "event_date", "v0", "v1", "v2", "token_span"] + (["name_span"] if fmt == "json" else [])
        p_data = pinned_array((total_bytes,), np.uint8)
        p_data[:] = data
        p_off = pinned_array((n + 1,), np.uint64)
        p_off[:] = rec_off
        p_cols = pinned_cols(n, e2e_cols)
        p_none = pinned_cols(n, [])
        p_status = pinned_array((n,), np.uint8)
        p_first = pinned_array((n + 1,), np.uint32)
        p_pairs = pinned_array((max_pairs,), np.uint64)
        got = {}

        def call_full():
            _, st, fr, rows, prs = ctx.decode_zone_test(fmt, p_data, p_off, p_cols, p_status, p_first, p_pairs)
            got["full"] = (rows, prs.shape[0])

        def call_alerts():
            _, st, fr, rows, prs = ctx.decode_zone_test(fmt, p_data, p_off, p_none, p_status, p_first, p_pairs)
            got["alerts"] = (rows, prs.shape[0])

        e2e_steps = max(1, args.steps)
        t_full = e2e_region(env, call_full, e2e_steps)
        assert got["full"] == (n_rows, n_pairs), (got, n_rows, n_pairs)
        t_alerts = e2e_region(env, call_alerts, e2e_steps)
        assert got["alerts"] == (n_rows, n_pairs), (got, n_rows, n_pairs)
        h2d = total_bytes + 8 * (n + 1)
        d2h = len(e2e_cols) * 8 * n_rows + n + 4 * (n + 1) + 8 * n_pairs
        try:
            link = pcie_probe(torch, env.dist, dev, env.world)
            link["floor_ms_per_step_this_rank"] = max(h2d / (link["duplex_h2d_GBps"] * 1e6),
                                                      d2h / (link["duplex_d2h_GBps"] * 1e6))
        except Exception as e:  # context only
            link = {"error": repr(e)[:120]}
        e2e = {"value": n * env.world * e2e_steps / t_full, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": 1e3 * t_full / e2e_steps,
               "mode": "full: " + ",".join(e2e_cols) + " + status + first_row + alert pairs",
               "alerts_only": {"value": n * env.world * e2e_steps / t_alerts, "h2d_bytes_per_step": h2d,
                               "d2h_bytes_per_step": n + 4 * (n + 1) + 8 * n_pairs,
                               "ms_per_step": 1e3 * t_alerts / e2e_steps,
                               "mode": "status + first_row + alert pairs (no decoded columns copied back)"},
               "link": link}
    if env.rank != 0:
        return None
    peak, peak_src = measured_peaks()
    # B_dec = payload + 8 per record (offsets) + 64 per decoded row
    bytes_s8d = total_bytes + 8 * n + 64 * n_rows
    bytes_moved = total_bytes + 8 * (n + 1) + 72 * n_rows + n + 4 * (n + 1)   # nine 8-byte columns + status + first_row
    kname = {"json": "json_template_kernel", "enriched": "enriched_decode_kernel<dense>", "pb": "pb_decode_dense_kernel<false>"}[fmt]
    ach = bytes_s8d / (ms_dec * 1e-3) / 1e9
    tr = profile_traffic(kname, n)
    ms_step = ms_all / args.steps
    line = {
        "value": value, "ms_per_step": ms_step,
        "config": {"events_per_gpu_per_step": n, "zones": int(len(zones[3])), "alerts_per_step": n_pairs,
                   "avg_payload_bytes": total_bytes / n,
                   "l2": "inputs (%.0f MB payload + %.0f MB columns) exceed the 126 MB L2; no flush needed"
                         % (total_bytes / 1e6, 72 * n / 1e6),
                   "sharding": "Kafka murmur2(deviceToken) % n_gpus; zone table NCCL-broadcast once"},
        "e2e": e2e,
        # per step: (JSON: template pass,) decode dense pass, ordered pass (returns at once on a dense batch), zone
        # count, scan, emit
        "gpu_launches": (6 if fmt == "json" else 5) * args.steps,
        "roofline": {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "traffic": tr["dram_bytes"] if tr else None, "traffic_source": tr, "peak_source": peak_src,
                     "algorithmic_bytes": bytes_s8d,
                     "algorithmic_bytes_rule": "SURVEY.md §8(d): payload + 8/record + 64/row",
                     "bytes_moved_expected": bytes_moved, "duration_ms": ms_dec,
                     "duration": "mean over the timed steps, CUDA events around the decode stage on the launch stream "
                                 "(JSON: template pass + dense pass over what it left; protobuf: dense pass; + the ordered pass that returns at once + one small memset)",
                     "step_frac_hbm": (bytes_s8d + 24 * n_rows + 8 * n_pairs) / (ms_step * 1e-3) / 1e9 / peak},
        "kernels": {"decode": {"ms": ms_dec, "events_per_s": n / (ms_dec * 1e-3)}, "zone_test": zb,
                    "encode_inbound_payload (N2, outside the timed step)": enc,
                    "decode_merge_with_presence (outside the timed step)": mrg},
    }
    return line, zones, (data, rec_off)


# ------------------------------------------------------------------------------------------------ c3
def run_c3(args, env):
    torch, sw, ctx, dev, stream = env.torch, env.sw, env.ctx, env.dev, env.stream
    from sitewhere_b200.api import pinned_array
    import psutil
    _, batches = build_inputs("c3", args.events, env.rank, env.world)
    nb = len(batches)
    per = batches[0][1].shape[0] - 1
    n_total = nb * per
    payload_bytes = sum(int(b[1][-1]) for b in batches)
    dbat = [to_dev(torch, dev, d, o) for d, o in batches]
    id_lo = torch.arange(per, dtype=torch.int64, device=dev)
    id_hi = [torch.full((per,), b, dtype=torch.int64, device=dev) for b in range(nb)]
    d_status = torch.empty(per, dtype=torch.uint8, device=dev)
    d_rows = torch.zeros(1, dtype=torch.int32, device=dev)
    ctx.state_clear()
    ctx.state_reserve(32_000_000, 1 << 20)     # 1 M devices x (1 location + 16 names + 8 alert types) keys
    seq = [0]
    stride = per << 12

    def step(timed):
        for b in range(nb):
            ctx.decode_merge_dev("pb", dbat[b][0], dbat[b][1], per, seq[0], id_hi[b], id_lo, d_status, d_rows)
            seq[0] += stride

    with torch.cuda.stream(stream):
        # the first pass inserts the keys; the timed steps fold the same events again with later sequence
        # numbers into the now populated table (every row still finds its key by bytes and wins its slot)
        torch.cuda.synchronize()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        step(False)
        b_.record(stream)
        torch.cuda.synchronize()
        first_pass_ms = a.elapsed_time(b_)
        keys, kbytes = ctx.state_size()
        ms = timed_region(env, step, args.steps, max(args.warmup - 1, 2))
        assert int(d_rows.item()) == per and int(d_status.max().item()) == 0
        keys2, _ = ctx.state_size()
        assert keys2 == keys, (keys, keys2)
        # the unfused pair on batch 0, for comparison (decode to columns, then merge from the columns)
        cols = sw.DeviceCols(per, dev, optional=False)
        d_first = torch.empty(per + 1, dtype=torch.int32, device=dev)

        def t_of(fn, reps=3):
            fn()
            torch.cuda.synchronize()
            x, y = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            x.record(stream)
            for _ in range(reps):
                fn()
            y.record(stream)
            torch.cuda.synchronize()
            return x.elapsed_time(y) / reps
        ms_dec = t_of(lambda: ctx.decode_pb_dev(dbat[0][0], dbat[0][1], per, cols, d_status, d_first, d_rows))

        def unfused_merge():
            ctx.merge_dev(dbat[0][0], dbat[0][1], per, cols, per, d_rows, seq[0], id_lo, id_lo)
            seq[0] += stride
        ms_mrg = t_of(unfused_merge)
        ctx.state_size()
    ms_all = max_over_ranks(env, ms)
    value = n_total * env.world * args.steps / (ms_all * 1e-3)

    # ---- end to end: swgpu_decode_merge per batch on pinned host buffers (payload + offsets + 16 B of event id per
    # record up, one status byte per record down; the state stays on the GPU)
    e2e = {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    if not args.no_e2e:
        need = payload_bytes + nb * (per + 1) * 8 + (nb + 1) * per * 8 + per
        avail = psutil.virtual_memory().available
        use = nb if need * 2 < avail else max(1, int(nb * avail / (need * 2.5)))
        pb = []
        for b in range(use):
            d, o = batches[b]
            pd_ = pinned_array((d.shape[0],), np.uint8)
            pd_[:] = d
            po = pinned_array((per + 1,), np.uint64)
            po[:] = o
            ph = pinned_array((per,), np.uint64)
            ph[:] = b
            pb.append((pd_, po, ph))
        p_lo = pinned_array((per,), np.uint64)
        p_lo[:] = np.arange(per, dtype=np.uint64)
        p_st = pinned_array((per,), np.uint8)

        def call():
            for b in range(use):
                rows = ctx.decode_merge("pb", pb[b][0], pb[b][1], seq[0], pb[b][2], p_lo, p_st)
                seq[0] += stride
                assert rows == per
        e2e_steps = max(1, min(args.steps, 3))
        t = e2e_region(env, call, e2e_steps)
        ev = use * per
        e2e = {"value": ev * env.world * e2e_steps / t, "unit": UNIT,
               "h2d_bytes_per_step": int(sum(int(batches[b][1][-1]) for b in range(use)) + use * ((per + 1) * 8 + 16 * per)),
               "d2h_bytes_per_step": use * per, "steps": e2e_steps, "ms_per_step": 1e3 * t / e2e_steps,
               "events_per_step": ev,
               "mode": "swgpu_decode_merge per %d-record batch: payload + offsets + per-record event ids up, status down" % per}
        if use < nb:
            e2e["note"] = "host memory allowed only %d of the %d batches to be pinned at once" % (use, nb)
    if env.rank != 0:
        return None
    peak, peak_src = measured_peaks()
    # decode B_dec = payload + 8/record (the fused kernel writes no columns: the 64/row term is
    # what it saves) and merge B_mrg = 41/row streamed + 32/row table RMW (+ 48 per NEW key: none in a timed step)
    ms_step = ms_all / args.steps
    bytes_s8d = payload_bytes + 8 * n_total + 73 * n_total
    bytes_moved = payload_bytes + 8 * n_total + 16 * n_total + n_total + 128 * n_total
    ach = bytes_s8d / (ms_step * 1e-3) / 1e9
    tr = profile_traffic("pb_decode_dense_kernel<true>", per)
    line = {
        "value": value, "ms_per_step": ms_step,
        "config": {"events_per_gpu_per_step": n_total, "batches_per_step": nb, "records_per_batch": per,
                   "devices": N_DEVICES, "state_keys": keys, "avg_payload_bytes": payload_bytes / n_total,
                   "state": "carried across steps: the first pass (first_pass_ms) inserts the %d keys, the timed steps fold "
                            "the same events with later sequence numbers into the populated table" % keys,
                   "first_pass_ms": first_pass_ms,
                   "l2": "payload %.1f GB and the state table (128 B per slot, at most half full) exceed the 126 MB L2; "
                         "no flush needed" % (payload_bytes / 1e9)},
        "e2e": e2e,
        "gpu_launches": 2 * nb * args.steps,   # fused decode+merge kernel + fix-up pass per batch
        "roofline": {"bound": "hbm", "kernel": "pb_decode_dense_kernel<true> (fused decode + merge)", "achieved": ach,
                     "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": tr["dram_bytes"] if tr else None,
                     "traffic_source": tr, "peak_source": peak_src, "algorithmic_bytes": bytes_s8d,
                     "algorithmic_bytes_rule": "SURVEY.md §8(d): decode payload + 8/record, merge 73/row (41 streamed + 32 RMW)",
                     "bytes_moved_expected": bytes_moved,
                     "bytes_moved_rule": "payload + offsets + 16 B ids + status + table: 64 B read, 64 B of dirty sectors per row",
                     "duration_ms": ms_step / nb,
                     "duration": "timed step / batches: the step is nothing but this kernel and its fix-up pass",
                     "merge_only_frac_hbm_unfused": 73 * per / (ms_mrg * 1e-3) / 1e9 / peak,
                     "random_access_ceiling": random_access_ceiling(ms_step / nb, ms_mrg, per)},
        "kernels": {"fused_decode_merge": {"ms_per_batch": ms_step / nb, "events_per_s": per / (ms_step / nb * 1e-3)},
                    "unfused_decode_to_columns": {"ms_per_batch": ms_dec},
                    "unfused_merge_from_columns": {"ms_per_batch": ms_mrg, "rows_per_s": per / (ms_mrg * 1e-3),
                                                   "GBps_s8d": 73 * per / (ms_mrg * 1e-3) / 1e9}},
    }
    return line, None, batches


# ------------------------------------------------------------------------------------------------ c5
def run_c5(args, env):
    torch, sw, ctx, dev, stream = env.torch, env.sw, env.ctx, env.dev, env.stream
    from sitewhere_b200.api import pinned_array, pinned_cols
    zones, (data, rec_off) = build_inputs("c5", args.events, env.rank, env.world)
    n = rec_off.shape[0] - 1
    t0 = time.perf_counter()
    if env.rank == 0:
        ctx.set_zones(*zones)
    t_table = time.perf_counter() - t0
    table_bytes = ctx.zone_table_bytes() if env.rank == 0 else 0
    if env.world > 1:
        from sitewhere_b200.shard import broadcast_zone_table
        broadcast_zone_table(ctx, env.dist, dev, env.rank)
    # cut into polls of at most 64 KB of payload
    bounds = [0]
    while bounds[-1] < n:
        a = bounds[-1]
        b = int(np.searchsorted(rec_off, rec_off[a] + np.uint64(C5_POLL_BYTES), side="right")) - 1
        bounds.append(min(max(b, a + 1), n))
    polls = list(zip(bounds[:-1], bounds[1:]))
    per = max(b - a for a, b in polls)
    # device-resident form: the whole stream in HBM, one decode + zone test per poll
    d_bytes, d_off = to_dev(torch, dev, data, rec_off)
    cols = sw.DeviceCols(per + 64, dev, optional=False)
    d_status = torch.empty(per + 64, dtype=torch.uint8, device=dev)
    d_first = torch.empty(per + 65, dtype=torch.int32, device=dev)
    d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
    d_pairs = torch.empty(4 * per + 1024, dtype=torch.int64, device=dev)
    d_npairs = torch.zeros(1, dtype=torch.int64, device=dev)
    off_views = [d_off[a:] for a, b in polls]

    def step(timed):
        for k, (a, b) in enumerate(polls):
            ctx.decode_zone_test_dev("pb", d_bytes, off_views[k], b - a, cols, d_status, d_first, d_nrows, d_pairs, d_npairs)

    with torch.cuda.stream(stream):
        ms = timed_region(env, step, args.steps, args.warmup)
    ms_all = max_over_ranks(env, ms)
    value = n * env.world * args.steps / (ms_all * 1e-3)
    # ---- the stream as a consumer sees it: one swgpu_decode_zone_test call per poll on pinned buffers
    p_data = pinned_array((int(rec_off[-1]),), np.uint8)
    p_data[:] = data
    p_off = pinned_array((n + 1,), np.uint64)
    p_off[:] = rec_off
    e2e_cols = ["meta", "event_date", "v0", "v1", "v2", "token_span"]
    p_cols = pinned_cols(per + 64, e2e_cols)
    p_status = pinned_array((per + 64,), np.uint8)
    p_first = pinned_array((per + 65,), np.uint32)
    p_pairs = pinned_array((4 * per + 1024,), np.uint64)
    lat = []
    tot_pairs = [0]

    def stream_pass(record):
        tot_pairs[0] = 0
        for a, b in polls:
            t1 = time.perf_counter()
            _, st, fr, rows, prs = ctx.decode_zone_test("pb", p_data, p_off[a: b + 1], p_cols, p_status, p_first, p_pairs)
            if record:
                lat.append(time.perf_counter() - t1)
            tot_pairs[0] += prs.shape[0]
    e2e_steps = max(1, args.steps)
    stream_pass(False)
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        stream_pass(True)
    t = max_over_ranks(env, time.perf_counter() - t0)
    lat_us = np.array(lat) * 1e6
    p50, p99, pmax = [max_over_ranks(env, float(x)) for x in
                      (np.percentile(lat_us, 50), np.percentile(lat_us, 99), lat_us.max())]
    # ---- several consumer threads per GPU, one context each (the zone table is copied between contexts as bytes):
    # what one GPU sustains when a single caller's round trip is not the limit
    T = 4
    thr = None
    try:
        nbytes = ctx.zone_table_bytes()
        blob = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        with torch.cuda.stream(stream):
            ctx.zone_table_export(blob)
        torch.cuda.synchronize()
        workers = []
        for k in range(T):
            c = sw.Context(env.local)
            c.zone_table_import(blob, nbytes)
            workers.append(dict(ctx=c, cols=pinned_cols(per + 64, e2e_cols), status=pinned_array((per + 64,), np.uint8),
                                first=pinned_array((per + 65,), np.uint32), pairs=pinned_array((4 * per + 1024,), np.uint64),
                                lat=[]))
        torch.cuda.synchronize()
        del blob

        def consume(w, k, record):
            for a, b in polls[k::T]:
                t1 = time.perf_counter()
                w["ctx"].decode_zone_test("pb", p_data, p_off[a: b + 1], w["cols"], w["status"], w["first"], w["pairs"])
                if record:
                    w["lat"].append(time.perf_counter() - t1)

        def all_threads(record):
            ts = [threading.Thread(target=consume, args=(w, k, record)) for k, w in enumerate(workers)]
            for t_ in ts:
                t_.start()
            for t_ in ts:
                t_.join()
        all_threads(False)
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            all_threads(True)
        tt = max_over_ranks(env, time.perf_counter() - t0)
        l4 = np.concatenate([np.array(w["lat"]) for w in workers]) * 1e6
        thr = {"threads": T, "value": n * env.world * e2e_steps / tt, "ms_per_step": 1e3 * tt / e2e_steps,
               "poll_latency_us": {"p50": max_over_ranks(env, float(np.percentile(l4, 50))),
                                   "p99": max_over_ranks(env, float(np.percentile(l4, 99)))},
               "note": "%d consumer threads per GPU, one context each, polls dealt round-robin" % T}
        for w in workers:
            w["ctx"].close()
    except Exception as e:  # context only
        thr = {"error": repr(e)[:160]}
    if env.rank != 0:
        return None
    peak, peak_src = measured_peaks()
    total_bytes = int(rec_off[-1])
    bytes_s8d = total_bytes + 8 * n + 64 * n
    ms_step = ms_all / args.steps
    h2d = total_bytes + 8 * (n + len(polls))
    d2h = len(e2e_cols) * 8 * n + n + 4 * (n + len(polls)) + 8 * tot_pairs[0]
    ach = bytes_s8d / (ms_step * 1e-3) / 1e9
    line = {
        "value": value, "ms_per_step": ms_step,
        "config": {"events_per_gpu_per_step": n, "polls_per_step": len(polls), "events_per_poll": n / len(polls),
                   "poll_payload_bytes": C5_POLL_BYTES, "devices": 10_000_000, "zones": int(len(zones[3])),
                   "zone_table_bytes": table_bytes, "zone_table_build_s": t_table, "alerts_per_step": tot_pairs[0],
                   "l2": "launch-latency bound (one clear and two launches per ~1,040-event poll), not bandwidth: the 64 KB poll and "
                         "its columns fit L2, the %.1f GB zone table does not" % (table_bytes / 1e9)},
        "e2e": {"value": n * env.world * e2e_steps / t, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "ms_per_step": 1e3 * t / e2e_steps,
                "poll_latency_us": {"p50": p50, "p99": p99, "max": pmax, "polls_timed": len(lat),
                                    "note": "wall time of one swgpu_decode_zone_test call on the consumer thread, max over ranks"},
                "mode": "one call per poll: upload, decode, zone test, download of rows + alerts (small-batch path)",
                "consumer_threads": thr},
        "gpu_launches": 2 * len(polls) * args.steps,   # small-batch path: the decoder's ordered pass + the single-launch zone kernel
        "roofline": {"bound": "hbm", "kernel": "pb_decode_dense_kernel<false> per poll", "achieved": ach, "peak": peak,
                     "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes": bytes_s8d, "duration_ms": ms_step / len(polls),
                     "duration": "timed step / polls (decode + zone test of one poll: launch latency dominates)"},
        "kernels": {"per_poll_us_device_resident": 1e3 * ms_step / len(polls)},
    }
    return line, zones, (data, rec_off)


# ------------------------------------------------------------------------------------------------ main (ours)
def run_ours(args):
    import torch
    import torch.distributed as dist

    import sitewhere_b200 as sw

    env = Env()
    env.torch, env.dist, env.sw = torch, dist, sw
    env.world = int(os.environ.get("WORLD_SIZE", "1"))
    env.rank = int(os.environ.get("RANK", "0"))
    env.local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU path)")
    torch.cuda.set_device(env.local)
    env.dev = torch.device("cuda", env.local)
    if env.world > 1:
        dist.init_process_group("nccl", device_id=env.dev)
    env.stream = torch.cuda.Stream(device=env.dev)
    env.ctx = sw.Context(env.local, stream=env.stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if env.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    env.barrier = barrier

    w = WORKLOADS[args.workload]
    if args.workload in ("c1", "c2", "n1"):
        res = run_decode_zone(args, env, args.workload)
    elif args.workload == "c3":
        res = run_c3(args, env)
    else:
        res = run_c5(args, env)
    if env.rank == 0:
        body, zones, inputs = res
        line = {"metric": w["metric"], "value": body.pop("value"), "unit": UNIT, "n_gpus": env.world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": body.pop("ms_per_step"), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64", "data": "synthetic"}
        cfg = {"workload": w["name"]}
        cfg.update(body.pop("config"))
        line["config"] = cfg
        line.update(body)
        line["clocks"] = env.clocks
        if env.world == 1 and not args.no_cpu:
            import oracle
            oracle.build()
            cores = os.cpu_count() or 1
            ev, t_cpu, desc = cpu_leg(args.workload, zones, inputs, cores, args.cpu_events or CPU_SAMPLE[args.workload])
            line["cpu_baseline"] = {"value": ev / t_cpu, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": desc + " (C oracle; JVM path unavailable on this box: no JDK)"}
        print(json.dumps(line))
    env.ctx.close()
    if env.world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--events", type=int, default=0, help="events per GPU per step (default: the workload's size)")
    ap.add_argument("--cpu-events", type=int, default=0, help="CPU baseline sample size")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--no-extras", action="store_true", help="(accepted for compatibility; the other workloads are "
                                                             "separate --workload runs now)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if not args.events:
        args.events = WORKLOADS[args.workload]["events"]
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
