#!/usr/bin/env python
"""bench.py — events/sec of the inbound hot path (protobuf decode + geofence zone test) on B200.

Workload (BASELINE.json configs[1], "C2"): 10 M protobuf DeviceLocation payloads per GPU, decoded as
ProtobufDeviceEventDecoder does and tested against 1,024 concave 64-vertex zones as
ZoneTestRuleProcessor does.  A step is one pass of the path over the batch.

  value     : events/s, inputs resident in HBM, CUDA events on the launch stream, max over ranks
  e2e       : events/s through the host-buffer C-ABI calls (H2D of payloads, D2H of rows + alerts)
  roofline  : achieved algorithmic GB/s of the dominant kernel vs the measured HBM peak
  cpu_baseline : the CPU oracle (restatement of the JVM path) on a bounded sample, all host cores

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--events E]
Multi-GPU: torchrun launches one rank per GPU; events are sharded by Kafka's partitioner on the
device token, the zone table is built on rank 0 and replicated with one NCCL broadcast; there is
no per-batch collective (weak scaling).
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

METRIC = "events/sec decode+zone-test"
UNIT = "events/s"
N_DEVICES = 1_000_000
ZONE_GRID = 32          # 32 x 32 = 1,024 zones
ZONE_VERTS = 64


WORKLOAD = "C2: 10M protobuf DeviceLocation decode + zone test vs 1,024 concave 64-vertex zones"

def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


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


# ------------------------------------------------------------------------------------------------ workload
def make_workload(n_events: int, rank: int, world: int):
    from sitewhere_b200 import synth as S

    zones = S.star_zones(ZONE_GRID, nvert=ZONE_VERTS, seed=S.SEED_BASE + 1)
    lon, lat = S.locations_for_zones(n_events, *zones[:3], seed=S.SEED_BASE + 1 + 1000 * rank)
    devices = S.devices_of_shard(N_DEVICES, rank, world)
    data, rec_off, truth = S.pb_locations(n_events, N_DEVICES, lon, lat, seed=S.SEED_BASE + 1 + 1000 * rank,
                                          devices=devices)
    return zones, data, rec_off, truth


def cpu_pass(zones, data, rec_off, threads: int):
    """The oracle's decode + index-free zone loop over a sample, split over `threads` host threads
    (ctypes releases the GIL).  Returns seconds."""
    import oracle
    from concurrent.futures import ThreadPoolExecutor

    n = rec_off.shape[0] - 1
    bounds = np.linspace(0, n, threads + 1).astype(np.int64)

    def work(k):
        a, b = int(bounds[k]), int(bounds[k + 1])
        if a == b:
            return 0
        off = rec_off[a: b + 1] - rec_off[a]
        chunk = data[int(rec_off[a]): int(rec_off[b])]
        cols, status, first_row, total = oracle.decode_pb(chunk, off, max_rows=b - a)
        pairs, tp = oracle.zone_test(*zones, cols.v1[:total], cols.v0[:total], meta=cols.meta[:total],
                                     max_pairs=max(total, 1))
        return tp

    oracle.lib()
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(threads)))
    return time.perf_counter() - t0


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The JVM path cannot run
    here (no JDK/jars), so this times the CPU oracle (a C restatement, kind 'port') on all host
    cores, on a bounded sample of the same workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    sample = args.cpu_events or 2_000_000
    zones, data, rec_off, _ = make_workload(sample, 0, 1)
    for _ in range(args.warmup):
        cpu_pass(zones, data[: int(rec_off[sample // 10])], rec_off[: sample // 10 + 1], cores)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_pass(zones, data, rec_off, cores)
    v = sample * args.steps / t
    line = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64", "data": "synthetic",
            "impl": "reference",
            "config": {"workload": WORKLOAD, "sample_events_per_step": sample, "zones": ZONE_GRID * ZONE_GRID,
                       "vertices": ZONE_VERTS, "devices": N_DEVICES},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{sample} events of the C2 workload per step (JVM path unavailable: no JDK on this box)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))



def pcie_probe(torch, dev, nbytes=256 << 20, reps=4):
    """Host<->device copy rates of this box (pinned memory): each direction alone, then both at once."""
    hin = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    hout = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    din = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dout = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(up, down):
        torch.cuda.synchronize()
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
        return (gb / ev[0].elapsed_time(ev[1]) if up else None, gb / ev[2].elapsed_time(ev[3]) if down else None)

    run(True, True)
    h, _ = run(True, False)
    _, d = run(False, True)
    dh, dd = run(True, True)
    return {"h2d_GBps": h, "d2h_GBps": d, "duplex_h2d_GBps": dh, "duplex_d2h_GBps": dd}

def run_extras(ctx, sw, torch, stream, dev, peak):
    """Secondary measurements of the other rows of the scope table (not the headline): the JSON decoder on
    the C1 shape and decode + device-state merge on the C3 shape, device-resident, CUDA events."""
    from sitewhere_b200 import synth as S
    out = {}

    def timeit(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream)
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    with torch.cuda.stream(stream):
        # ---- C1 shape: JSON DeviceMeasurement requests (250 K distinct payloads tiled to 2 M records)
        data, rec_off, _ = S.json_measurements(250_000, 100_000, seed=S.SEED_BASE)
        reps = 8
        lens = np.diff(rec_off.astype(np.int64))
        data = np.tile(data, reps)
        rec_off = np.concatenate([[0], np.cumsum(np.tile(lens, reps))]).astype(np.uint64)
        n = rec_off.shape[0] - 1
        d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
        cols = sw.DeviceCols(n, dev, optional=False)
        d_status = torch.empty(n, dtype=torch.uint8, device=dev)
        d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
        ms = timeit(lambda: ctx.decode_json_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows))
        assert int(d_nrows.item()) == n and int(d_status.max().item()) == 0
        nbytes = data.shape[0] + 8 * (n + 1) + 72 * n + n + 4 * (n + 1)
        out["json_decode"] = {"workload": "C1 shape: %d JSON DeviceMeasurement requests, %.0f B avg" % (n, data.shape[0] / n),
                              "ms": ms, "events_per_s": n / (ms * 1e-3), "algorithmic_bytes": nbytes,
                              "GBps": nbytes / (ms * 1e-3) / 1e9, "frac_hbm": nbytes / (ms * 1e-3) / 1e9 / peak}
        del d_bytes, d_off, cols, d_status, d_first
        # ---- C3 shape (1/10 scale): 10 M mixed protobuf events, 1 M devices: decode, then merge into device state
        n = 10_000_000
        data, rec_off, _ = S.pb_mixed(n, 1_000_000, seed=S.SEED_BASE + 2)
        d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
        cols = sw.DeviceCols(n, dev, optional=False)
        d_status = torch.empty(n, dtype=torch.uint8, device=dev)
        d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
        ids = torch.arange(n, dtype=torch.int64, device=dev)
        ctx.state_clear()
        ctx.state_reserve(20_000_000, 20_000_000 * 40)
        ms_dec = timeit(lambda: ctx.decode_pb_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows))
        seq = [0]

        def merge():
            ctx.merge_dev(d_bytes, d_off, n, cols, n, d_nrows, seq[0], ids, ids)
            seq[0] += n
        ms_mrg = timeit(merge, reps=3)
        keys, kbytes = ctx.state_size()
        bytes_mrg = 41 * n + 64 * n          # streamed columns + one 64-byte slot per row
        out["merge"] = {"workload": "C3 shape at 1/10 scale: %d mixed protobuf events, 1M devices, %d state keys" % (n, keys),
                        "decode_ms": ms_dec, "merge_ms": ms_mrg, "events_per_s_decode_plus_merge": n / ((ms_dec + ms_mrg) * 1e-3),
                        "merge_events_per_s": n / (ms_mrg * 1e-3), "algorithmic_bytes": bytes_mrg,
                        "GBps": bytes_mrg / (ms_mrg * 1e-3) / 1e9, "frac_hbm": bytes_mrg / (ms_mrg * 1e-3) / 1e9 / peak}
        ctx.state_clear()
        # ---- the same batch end to end through the host-buffer call of the state consumer (pinned payload in,
        #      state on the GPU; one upload of payload + offsets + ids, a 10 MB status download)
        try:
            from sitewhere_b200.api import pinned_array
            p_data = pinned_array((data.shape[0],), np.uint8)
            p_data[:] = data
            p_off = pinned_array((n + 1,), np.uint64)
            p_off[:] = rec_off
            p_ids = pinned_array((n,), np.uint64)
            p_ids[:] = np.arange(n, dtype=np.uint64)
            p_st = pinned_array((n,), np.uint8)
            ctx.state_reserve(20_000_000, 20_000_000 * 40)
            ctx.decode_merge("pb", p_data, p_off, 0, p_ids, p_ids, p_st)
            t0 = time.perf_counter()
            for k in range(3):
                ctx.decode_merge("pb", p_data, p_off, (k + 1) * n, p_ids, p_ids, p_st)
            dt = (time.perf_counter() - t0) / 3
            out["merge"]["e2e_decode_merge_events_per_s"] = n / dt
            out["merge"]["e2e_h2d_bytes"] = int(data.shape[0] + 8 * (n + 1) + 16 * n)
            ctx.state_clear()
        except Exception as e:  # secondary number
            out["merge"]["e2e_error"] = repr(e)[:160]
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    import sitewhere_b200 as sw
    from sitewhere_b200 import _lib as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.events
    zones, data, rec_off, truth = make_workload(n, rank, world)

    stream = torch.cuda.Stream(device=dev)
    ctx = sw.Context(local, stream=stream.cuda_stream)

    # ---- zone table: built on rank 0, replicated by one NCCL broadcast (configuration change only)
    if rank == 0:
        ctx.set_zones(*zones)
    if world > 1:
        from sitewhere_b200.shard import broadcast_zone_table
        broadcast_zone_table(ctx, dist, dev, rank)

    # ---- device-resident inputs and outputs
    total_bytes = int(rec_off[-1])
    h_bytes = torch.from_numpy(data).pin_memory()
    h_off = torch.from_numpy(rec_off.view(np.int64)).pin_memory()
    d_bytes = torch.empty(total_bytes + 64, dtype=torch.uint8, device=dev)
    d_bytes[:total_bytes].copy_(h_bytes)
    d_off = h_off.to(dev)
    cols = sw.DeviceCols(n, dev, optional=False)
    d_status = torch.empty(n, dtype=torch.uint8, device=dev)
    d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
    max_pairs = n
    d_pairs = torch.empty(max_pairs, dtype=torch.int64, device=dev)
    d_npairs = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()

    def step():
        ctx.decode_pb_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        ctx.zone_test_dev(cols.v1, cols.v0, cols.meta, n, d_nrows, d_pairs, d_npairs)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            step()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        n_rows = int(d_nrows.item())
        n_pairs = int(d_npairs.item())
        assert n_rows == n, (n_rows, n)
        assert int(d_status.max().item()) == 0

        # ---- per-kernel durations (each kernel alone, same inputs, inputs > L2)
        def time_kernel(fn, reps):
            fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(reps):
                fn()
            b.record(stream)
            torch.cuda.synchronize()
            return a.elapsed_time(b) / reps

        reps = max(3, args.steps)
        ms_dec = time_kernel(lambda: ctx.decode_pb_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows), reps)
        ms_zone = time_kernel(lambda: ctx.zone_test_dev(cols.v1, cols.v0, cols.meta, n, d_nrows, d_pairs, d_npairs), reps)
        # the timed region lasts milliseconds, nvidia-smi answers in ~100 ms: keep the same steps running for
        # about a second more so that the clock / throttle-reason samples are taken under the same load
        t_probe = time.perf_counter() + 1.2
        while time.perf_counter() < t_probe:
            for _ in range(20):
                step()
            torch.cuda.synchronize()
        clocks = sampler.stop() if rank == 0 else None

    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_all = float(tmax.item())
    value = n * world * args.steps / (ms_all * 1e-3)

    # ---- end to end through the host-buffer C ABI: H2D payloads + offsets, D2H rows + alerts
    # One call of swgpu_decode_zone_test per step on PINNED host buffers (what the Java binding's staging
    # segments are).  Copied back per event: meta, eventDate, lat, lon, elevation, token span (everything
    # a DeviceLocationCreateRequest + token holds; hashes are derivable, name columns do not apply to
    # locations), status, first_row, and the alert pairs.
    from sitewhere_b200.api import pinned_array, pinned_cols
    e2e_cols = ["meta", "event_date", "v0", "v1", "v2", "token_span"]
    e2e_value, h2d, d2h, e2e_steps, pcie = None, 0, 0, 0, None
    if not args.no_e2e:
        p_data = pinned_array((total_bytes,), np.uint8)
        p_data[:] = data
        p_off = pinned_array((n + 1,), np.uint64)
        p_off[:] = rec_off
        p_cols = pinned_cols(n, e2e_cols)
        p_status = pinned_array((n,), np.uint8)
        p_first = pinned_array((n + 1,), np.uint32)
        p_pairs = pinned_array((max_pairs,), np.uint64)
        e2e_steps = max(1, args.steps)

        def e2e_step():
            _, st, fr, rows, prs = ctx.decode_zone_test("pb", p_data, p_off, p_cols, p_status, p_first, p_pairs)
            return rows, prs.shape[0]

        for _ in range(2):
            r_rows, r_pairs = e2e_step()
        assert r_rows == n_rows and r_pairs == n_pairs, (r_rows, n_rows, r_pairs, n_pairs)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            r_rows, r_pairs = e2e_step()
        torch.cuda.synchronize()
        t_e2e = time.perf_counter() - t0
        te = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_value = n * world * e2e_steps / float(te.item())
        h2d = total_bytes + 8 * (n + 1)
        d2h = len(e2e_cols) * 8 * r_rows + n + 4 * (n + 1) + 8 * r_pairs
        # the link's own ceiling on this box, for context: pinned copies, each direction alone and both at once
        pcie = None
        if rank == 0:
            try:
                pcie = pcie_probe(torch, dev)
                pcie["floor_ms"] = max(h2d / (pcie["duplex_h2d_GBps"] * 1e6), d2h / (pcie["duplex_d2h_GBps"] * 1e6))
            except Exception as e:  # context only
                pcie = {"error": repr(e)[:120]}

    if rank == 0:
        peak, peak_src = measured_peaks()
        bytes_dec = total_bytes + 8 * (n + 1) + 72 * n_rows + n + 4 * (n + 1)
        bytes_zone = 24 * n + 8 * n_pairs
        gbs_dec = bytes_dec / (ms_dec * 1e-3) / 1e9
        gbs_zone = bytes_zone / (ms_zone * 1e-3) / 1e9
        # The decode stage is pb_decode_dense_kernel plus the ordered pass that returns at once (~3 us); the zone
        # stage is three launches of which zone_count_kernel is the largest (84 % of the stage in the committed
        # launch list, profiles/r01d_launches.csv).  Dominant kernel = the longer single launch.
        ZONE_COUNT_SHARE = 0.843
        dom = "pb_decode_kernel" if ms_dec >= ZONE_COUNT_SHARE * ms_zone else "zone_test (zone_count_kernel + zone_scan_mc_kernel + zone_emit_kernel)"
        ach = gbs_dec if dom == "pb_decode_kernel" else gbs_zone
        # DRAM bytes per launch from the committed `ncu --set full` capture of this very workload
        # (profiles/r01d_ncu_full_summary.md): decode 710.3 + 733.3 MB; zone 258.6 + 67.1 (count) + 0.3 (scan) + 80.4 + 7.4 (emit) MB
        traffic = None
        if n == 10_000_000:
            traffic = 1.4435e9 if dom == "pb_decode_kernel" else 413.8e6
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_all / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "events_per_gpu_per_step": n, "zones": ZONE_GRID * ZONE_GRID, "vertices": ZONE_VERTS,
                       "devices": N_DEVICES, "alerts_per_step": n_pairs,
                       "l2": "inputs (%.0f MB payload + %.0f MB columns) exceed the 126 MB L2; no flush needed"
                             % (total_bytes / 1e6, 72 * n / 1e6),
                       "sharding": "Kafka murmur2(deviceToken) % n_gpus; zone table NCCL-broadcast once"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": (n * world / e2e_value * 1e3) if e2e_value else None,
                    "pcie": pcie},
            # per step: pb_decode_kernel<dense>, pb_decode_kernel<ordered> (returns at once on a dense batch),
            # zone_count_kernel, zone_scan_mc_kernel, zone_emit_kernel
            "gpu_launches": 5 * args.steps,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                         "note": "pb_decode_dense_kernel is the longest launch of the step and is HBM-bound; the zone test "
                                 "(kernels.zone_test_kernel) is bound by instruction issue at ~17 active lanes per "
                                 "warp instruction on an L2-resident table, not by HBM"},
            "kernels": {"pb_decode_kernel": {"ms": ms_dec, "algorithmic_bytes": bytes_dec, "GBps": gbs_dec,
                                             "frac_hbm": gbs_dec / peak, "events_per_s": n / (ms_dec * 1e-3)},
                        "zone_test_kernel": {"ms": ms_zone, "algorithmic_bytes": bytes_zone, "GBps": gbs_zone,
                                             "frac_hbm": gbs_zone / peak, "bound": "instruction issue / divergence (zone table in L2)",
                                             # from the committed ncu capture of this workload (profiles/r01d_ncu_full_summary.md)
                                             "ncu": {"zone_count_kernel": {"fp64_pipe_pct": 23.6, "issue_active_pct": 56.0,
                                                                           "active_lanes_per_inst": 17.3}},
                                             "locations_per_s": n / (ms_zone * 1e-3)}},
            "clocks": clocks,
        }
        if world == 1 and not args.no_extras:
            try:
                line["extras"] = run_extras(ctx, sw, torch, stream, dev, peak)
            except Exception as e:  # secondary numbers must never cost the headline line
                line["extras"] = {"error": repr(e)[:200]}
        # ---- CPU baseline on a bounded sample (rank 0, N = 1 only)
        if world == 1 and not args.no_cpu:
            import oracle
            oracle.build()
            cores = os.cpu_count() or 1
            sample = min(n, args.cpu_events or 4_000_000)
            t_cpu = cpu_pass(zones, data[: int(rec_off[sample])], rec_off[: sample + 1], cores)
            line["cpu_baseline"] = {"value": sample / t_cpu, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"first {sample} events of the same batch, decode + index-free zone loop "
                                              f"(C oracle; JVM path unavailable on this box: no JDK)"}
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--events", type=int, default=10_000_000, help="events per GPU per step")
    ap.add_argument("--cpu-events", type=int, default=0, help="CPU baseline sample size")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary JSON / merge measurements")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
