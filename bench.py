#!/usr/bin/env python
"""bench.py — headline benchmark of the index-organized read path on B200.

Workload (BASELINE.json configs[1]): YCSB int table (4-byte keys, 1000-byte payload), 10 M tuples per
GPU, 10 % of rows overwritten by committed updates (version chains), scan_mode range scans of 1000
rows with the MVCC visibility filter (IndexScanExecutor range semantics) at a read timestamp in the
middle of the update commit ids. A "step" = one batch of `--scans` scans through the scan kernels.
Secondary leg in the same JSON line ("probe"): batches of 2^20 uniform point probes on the same table.

The GPU legs use product code only: the table is generated on the device with the reference's row
function in the reference's key order (stage_gpu_snapshot_generate_ycsb) and the committed updates are
installed by the refresh kernels (stage_gpu_snapshot_apply_updates); key streams come from the package's
restatement of the drivers' generators. oracle/ is touched only by the cpu_baseline leg (and by
--impl reference), which builds the same table with the reference's own loader, times the reference's
CPU scan path on it, and — having a reference-format tree at hand — also times the record-layout
transform on that tree ("transform") and cross-checks a sample of scans and probes ("parity_check").
`--table host` runs the GPU legs on the flattened host tree instead.

  value     rows/s, whole job (all ranks), inputs and outputs resident in HBM, CUDA-event timed
  e2e       the same metric through the host-buffer C ABI call (pinned host start keys in, pinned host
            rows out; the H2D/D2H copies are inside the timed region)
  roofline  algorithmic bytes of the dominant kernel / its CUDA-event duration vs the measured HBM copy peak
  cpu_baseline  the reference's own CPU path (oracle/_ref when its prebuilt library is present, else
            the C port) on this box's host cores over a bounded sample of the same scans

`--impl reference` times that CPU path alone, with every host thread, on the same metric/config.

N > 1 (torchrun, one rank per GPU): the table is range-partitioned in reference key order; every rank
holds --rows-multi tuples (default 125 M = BASELINE configs[4], 1 B tuples over 8 GPUs; weak scaling: the
per-GPU step is the same 4096 x 1000-row scans) built on the device plus a halo of the next shard's first
rows, with 10 % of its rows updated; scans run shard-local (no data-path collective); probes are routed
to their owning shard by kernels that store keys / result rows straight into the peers' memory over
NVLink (csrc/exchange.cuh), with the NCCL all-to-all form timed beside it.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K_BYTES, P_BYTES = 4, 1000
ROW_ALGO_BYTES = 2 * (K_BYTES + P_BYTES + 4)       # record read + result write per emitted row
PROBE_HIT_BYTES = K_BYTES + 2 * (K_BYTES + P_BYTES + 4)
PROBE_MISS_BYTES = K_BYTES + 1
# dram__bytes_read.sum + dram__bytes_write.sum of one gather_rows_kernel launch of the default workload, from the
# committed `ncu --set full` capture (profiles/); None until a capture of the current kernel exists.
NCU_TRAFFIC_PER_LAUNCH = 7.769e9   # profiles/r1_v10_ncu_summary.txt: 3.925 GB read + 3.844 GB written


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000, help="tuples per GPU at N = 1 (BASELINE configs[1])")
    ap.add_argument("--rows-multi", type=int, default=125_000_000,
                    help="tuples per GPU at N > 1 (BASELINE configs[4]: 1 B tuples over 8 GPUs)")
    ap.add_argument("--read-ts", choices=["first", "mid", "latest"], default="mid",
                    help="read id of the timed scans / probes relative to the update commit ids (SURVEY 8d config 2)")
    ap.add_argument("--table-scan", action="store_true",
                    help="scans without the visibility filter (TableScanExecutor semantics) instead of IndexScanExecutor's")
    ap.add_argument("--routed-scans", type=int, default=1024,
                    help="N > 1: scans per GPU per step of the routed-scan leg (start keys over the global table); 0 = skip")
    ap.add_argument("--a2a-steps", type=int, default=2, help="steps of the NCCL all-to-all comparison leg (N > 1)")
    ap.add_argument("--scans", type=int, default=4096, help="scans per step per GPU")
    ap.add_argument("--scan-size", type=int, default=1000)
    ap.add_argument("--probes", type=int, default=1 << 20, help="probes per probe step per GPU")
    ap.add_argument("--e2e-scans", type=int, default=1024)
    ap.add_argument("--cpu-scans", type=int, default=0, help="scans in the CPU-baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="scan", choices=["scan", "config1", "string", "tpcc"],
                    help="'scan' = BASELINE configs[1] (the headline, default); 'string' = configs[2] (100 M string keys, zipf "
                         "0.9 probes); 'tpcc' = configs[3] (w=30 stock-level / Q2 read pattern). The extra workloads are "
                         "single-GPU legs for profiles/README.md, not the driver's metric.")
    ap.add_argument("--table", default="device", choices=["device", "host"],
                    help="N = 1: 'device' builds the table with the on-device generator + refresh kernels (no oracle in "
                         "the GPU leg); 'host' flattens a tree built by the reference loader (transform kernel)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self.stop = [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80}
        while not self.stop:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def __enter__(self):
        if self.nv:
            self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        if self.nv:
            self.t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ----------------------------------------------------------------------------------------------
# input table (host tree in the reference's byte format) and CPU baseline — the only users of oracle/
# ----------------------------------------------------------------------------------------------
def build_host_table(rows, updates):
    """The YCSB table exactly as the reference's loader builds it (single loader thread), then
    `updates` committed single-row updates with commit ids 10..10+updates-1 (keys uniform, seed 2)."""
    from oracle import pyoracle
    pyoracle.build_oracle()
    if pyoracle.have_ref():
        t, kind = pyoracle.RefTable(P_BYTES, leaf_pool_blocks=max(rows // 20, 4096)), "reference"
    else:
        t, kind = pyoracle.OracleTable(P_BYTES), "port"
    assert t.load_ycsb_int(0, rows) == rows
    if updates:
        uk = update_keys(rows, updates).view(np.uint8).reshape(-1, 4)
        assert t.commit_update_batch(uk, FIRST_CID) == updates
    return t, kind


def cpu_scan_sample(table, start_keys, scan_size, read_ts, threads, vis=1):
    sec, rows = table.time_scan(start_keys, scan_size, read_ts, vis, threads, pin=True)
    return rows / sec, rows, sec


def calibrated_scan_count(table, rows, scan_size, read_ts, threads, target_s):
    """Number of scans that keeps one CPU sample near target_s seconds (pilot run of 64 scans/thread)."""
    pilot = scan_start_keys(rows, 64 * threads, seed=7).view(np.uint8).reshape(-1, 4)
    sec, _ = table.time_scan(pilot, scan_size, read_ts, 1, threads, pin=True)
    per_scan = max(sec / len(pilot), 1e-9)
    return int(min(max(target_s / per_scan, 64 * threads), 4_000_000))


def scan_start_keys(rows, n, seed):
    from stage_indexorganized_b200 import ycsb
    return (ycsb.uniform_keys(rows, n, seed=seed) - 1).astype(np.uint32)


FIRST_CID = 10


def update_keys(rows, updates):
    """Keys of the committed updates: uniform over the existing keys 0..rows-1 (FastRandom seed 2); update i commits
    with id FIRST_CID + i and rewrites one 100-byte column with (uint8) commit id."""
    from stage_indexorganized_b200 import ycsb
    return (ycsb.uniform_keys(rows, updates, seed=2) - 1).astype(np.uint32)


def apply_updates_on_device(snap, keys_u32, torch, dev):
    """The same committed updates, installed on the GPU snapshot (stage_gpu_snapshot_apply_updates)."""
    n = len(keys_u32)
    if n == 0:
        return 0
    d_keys = torch.from_numpy(np.ascontiguousarray(keys_u32, np.uint32).view(np.uint8).reshape(-1, 4)).to(dev)
    cids = torch.arange(FIRST_CID, FIRST_CID + n, dtype=torch.int32, device=dev)
    patches = (cids & 0xFF).to(torch.uint8)[:, None].expand(n, 100).contiguous()
    return snap.apply_updates(d_keys, patches, cids, patch_off=100, device=True)


# ----------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path, all host threads, bounded sample per step."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    rows = args.rows
    updates = rows // 10
    read_ts = {"first": FIRST_CID, "mid": FIRST_CID + updates // 2, "latest": FIRST_CID + updates}[args.read_ts]
    vis = 0 if args.table_scan else 1
    table, kind = build_host_table(rows, updates)
    # each step is a bounded sample: ~4 s of work on all host threads
    n = args.cpu_scans or calibrated_scan_count(table, rows, args.scan_size, read_ts, threads, 4.0)
    times, total_rows = [], 0
    for step in range(args.warmup + args.steps):
        sk = scan_start_keys(rows, n, seed=100 + step).view(np.uint8).reshape(-1, 4)
        rps, r, sec = cpu_scan_sample(table, sk, args.scan_size, read_ts, threads, vis)
        if step >= args.warmup:
            times.append(sec)
            total_rows += r
    value = total_rows / sum(times)
    line = {
        "impl": "reference", "metric": "YCSB scan rows/s (1000-row MVCC range scans, 10M-tuple int table)",
        "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, world=1, table="reference loader (host tree)"),
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": threads, "kind": kind,
                         "sample": f"{n} scans x {args.scan_size} rows per step, {args.steps} steps, {threads} pinned threads"},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world, table):
    return {"workload": "BASELINE configs[1]: YCSB scan_mode (-n) range scan with MVCC visibility, int keys, "
                        f"{args.rows} tuples per GPU, scan_size {args.scan_size}",
            "rows_per_gpu": args.rows, "payload_bytes": P_BYTES, "key_bytes": K_BYTES,
            "scans_per_step_per_gpu": args.scans, "scan_size": args.scan_size,
            "probes_per_step_per_gpu": args.probes, "updates_per_gpu": args.rows // 10,
            "read_ts": {"first": "first update commit id", "mid": "middle of the update commit ids",
                        "latest": "last update commit id"}[args.read_ts],
            "visibility": "none (TableScanExecutor)" if args.table_scan else "IndexScanExecutor range rule", "table": table,
            "partitioning": f"range x{world}" if world > 1 else "single shard",
            "l2": "inputs larger than L2 (table >= 10 GB, 4 GB of rows written per step); no explicit flush"}


def _time_steps(torch, fn, warmup, steps):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(warmup + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def run_extra_workload(args):
    """Single-GPU legs for the other BASELINE configs (device-resident inputs and outputs, CUDA-event timed)."""
    import torch
    import stage_indexorganized_b200 as sg
    from stage_indexorganized_b200 import tpcc, ycsb
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    sg.init(0)
    peak, peak_src = peaks()
    nb = args.warmup + args.steps
    if args.workload == "config1":
        # BASELINE configs[0]: ./ycsb -k 10000 -b 4 -o 10 -u 0 -z 0.2 — read-only point probes, int keys, 10 000 rows.
        # (The reference's own CLI aborts at HEAD with an out-of-range key pick, SURVEY.md §3.5, so the CPU side is the
        # reference's BTree::Read driven by the harness over the same key stream, at 4 threads like -b 4 and at all.)
        n_rows, n = 10_000, args.probes
        snap = sg.Snapshot.generate_ycsb(n_rows, P_BYTES)
        snap.set_stream(torch.cuda.current_stream().cuda_stream)
        z = ycsb.Zipf(n_rows - 1, 0.2, seed=1).draw(n * nb).astype(np.uint32).reshape(nb, n)
        d_keys = torch.from_numpy(z.view(np.int32)).to(dev)
        found = torch.zeros((n,), dtype=torch.uint8, device=dev)
        ts = torch.zeros((n,), dtype=torch.int32, device=dev)
        kout = torch.empty((n,), dtype=torch.int32, device=dev)
        pay = torch.empty((n, P_BYTES), dtype=torch.uint8, device=dev)
        ms = _time_steps(torch, lambda b: snap.probe_device(d_keys[b], n, 1, found, ts, kout, pay), args.warmup, args.steps)
        hits = int(found.sum().item())
        algo = hits * PROBE_HIT_BYTES + (n - hits) * PROBE_MISS_BYTES
        cpu = None
        if not args.no_cpu_baseline:
            from oracle import pyoracle
            pyoracle.build_oracle()
            t = pyoracle.RefTable(P_BYTES) if pyoracle.have_ref() else pyoracle.OracleTable(P_BYTES)
            assert t.load_ycsb_int(0, n_rows) == n_rows
            keys = np.ascontiguousarray(z.reshape(-1)[: 1 << 22]).view(np.uint8).reshape(-1, 4)
            ok = all(np.array_equal(g, w) for g, w in zip(snap.probe(keys[:65536], 1), t.probe(keys[:65536], 1)))
            cpu = {"kind": "reference" if pyoracle.have_ref() else "port", "unit": "probes/s", "parity_sample_ok": bool(ok)}
            for th in (4, os.cpu_count() or 1):
                sec, h = t.time_probe(keys, 1, th, pin=True)
                cpu[f"threads_{th}"] = len(keys) / sec
        line = {"metric": "YCSB read-only point probes/s (10k-tuple int table, zipf 0.2)", "value": n / (ms * 1e-3),
                "unit": "probes/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "dtype": "u8",
                "data": "synthetic", "config": {"workload": "BASELINE configs[0]: ./ycsb -k 10000 -b 4 -o 10 -u 0 -z 0.2",
                                                "probes_per_step": n, "hit_rate": hits / n,
                                                "note": "the 10 MB table lives in L2: the bound here is L2/SM throughput, not HBM"},
                "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": algo / (ms * 1e-3) / 1e9 / peak, "peak_source": peak_src},
                "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
        snap.close()
        return
    if args.workload == "string":
        n_rows, n = 100_000_000, args.probes
        snap, keys, _ = ycsb.build_string_table_device(n_rows)
        snap.set_stream(torch.cuda.current_stream().cuda_stream)
        z = np.clip(ycsb.Zipf(n_rows - 1, 0.9, seed=1).draw(n * nb).astype(np.int64), 1, n_rows) - 1
        d_keys = keys[torch.from_numpy(z).to(dev)].reshape(nb, n, 16).contiguous()
        found = torch.zeros((n,), dtype=torch.uint8, device=dev)
        ts = torch.zeros((n,), dtype=torch.int32, device=dev)
        kout = torch.empty((n, 16), dtype=torch.uint8, device=dev)
        pay = torch.empty((n, P_BYTES), dtype=torch.uint8, device=dev)
        ms = _time_steps(torch, lambda b: snap.probe_device(d_keys[b], n, 1, found, ts, kout, pay), args.warmup, args.steps)
        hits = int(found.sum().item())
        algo = hits * (16 + 2 * (16 + P_BYTES + 4)) + (n - hits) * 17
        line = {"metric": "YCSB string-key zipf-0.9 probes/s (100M-tuple table)", "value": n / (ms * 1e-3), "unit": "probes/s",
                "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "dtype": "u8", "data": "synthetic",
                "config": {"workload": "BASELINE configs[2]: YCSB string-key (-m) zipf 0.9 batched probes, 100000000 tuples, "
                                       "built on device (FastRandom(1804289383) keys, bulk load)", "probes_per_step": n,
                           "distinct_keys_per_step": int(np.unique(z[:n]).size), "hit_rate": hits / n},
                "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": algo / (ms * 1e-3) / 1e9 / peak, "peak_source": peak_src,
                             "note": "whole step (search + rows); zipf 0.9 repeats hot rows, so part of the row traffic is L2 hits"}}
        print(json.dumps(line), flush=True)
        snap.close()
        return
    # ---- tpcc: stock-level order_line scans + stock probes, Q2-shaped probe batches with S_QUANTITY push-down ------
    W = 30
    sk, sp = tpcc.stock_rows(W)
    ok_, op, _ = tpcc.order_line_rows(W)

    def bulk(keys, pays, K, P):
        order = np.lexsort(keys.T[::-1])                      # memcmp order of the raw key bytes (keys >= 16 B)
        snap, raw, pay, _ = sg.Snapshot.alloc_sorted(len(keys), K, P, rows_per_leaf=max(8, min(500, 60000 // (K + P + 24))))
        raw.copy_(torch.from_numpy(keys[order]))
        pay.copy_(torch.from_numpy(pays[order]))
        snap.seal_sorted()
        snap.set_stream(torch.cuda.current_stream().cuda_stream)
        return snap
    stock, ol = bulk(sk, sp, tpcc.STOCK_KEY, tpcc.STOCK_PAYLOAD), bulk(ok_, op, tpcc.OL_KEY, tpcc.OL_PAYLOAD)
    txns = 50_000                                             # 1 M order_line scans of 10 rows per step
    starts = [torch.from_numpy(tpcc.stock_level_requests(W, txns, seed=10 + b)[0]).to(dev) for b in range(nb)]
    ns = txns * 20
    sizes = torch.full((ns,), 10, dtype=torch.int32, device=dev)
    counts = torch.zeros((ns,), dtype=torch.int32, device=dev)
    kout = torch.empty((ns * 10, tpcc.OL_KEY), dtype=torch.uint8, device=dev)
    tso = torch.empty((ns * 10,), dtype=torch.int32, device=dev)
    pout = torch.empty((ns * 10, tpcc.OL_PAYLOAD), dtype=torch.uint8, device=dev)
    ms_scan = _time_steps(torch, lambda b: ol.scan_device(starts[b], sizes, ns, 10, 1, 1, counts, kout, tso, pout),
                          args.warmup, args.steps)
    rows = int(counts.sum().item())
    ol.enable_timing(True)
    ol.scan_device(starts[0], sizes, ns, 10, 1, 1, counts, kout, tso, pout)
    ol_plan_ms, ol_gather_ms = ol.last_timing()
    ol.enable_timing(False)
    rng = np.random.default_rng(3)
    n = args.probes
    pk = [torch.from_numpy(tpcc.stock_key(rng.integers(1, W + 1, n), rng.integers(1, tpcc.ITEMS + 1, n))).to(dev) for _ in range(nb)]
    found = torch.zeros((n,), dtype=torch.uint8, device=dev)
    spay = torch.empty((n, tpcc.STOCK_PAYLOAD), dtype=torch.uint8, device=dev)
    qty = torch.empty((n, 4), dtype=torch.uint8, device=dev)
    ms_full = _time_steps(torch, lambda b: stock.probe_device(pk[b], n, 1, found, None, None, spay), args.warmup, args.steps)
    ms_col = _time_steps(torch, lambda b: stock.probe_column_device(pk[b], n, 1, 0, 4, qty, found), args.warmup, args.steps)
    agg = [None]

    def agg_step(b):
        agg[0] = stock.probe_aggregate_i32(pk[b], 1, 0, 15, n=n)
    ms_agg = _time_steps(torch, agg_step, args.warmup, args.steps)
    ol_row, st_row = 2 * (tpcc.OL_KEY + tpcc.OL_PAYLOAD + 4), tpcc.STOCK_KEY + 2 * (tpcc.STOCK_KEY + tpcc.STOCK_PAYLOAD + 4)
    line = {"metric": "TPC-C w=30 stock-level / Q2 read pattern", "value": rows / (ms_scan * 1e-3), "unit": "order_line rows/s",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_scan, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "BASELINE configs[3]: TPC-C w=30, stock 3 M rows (16/400), order_line "
                                   f"{len(ok_)} rows (32/60), seeded row functions, bulk-loaded on device",
                       "order_line_scans_per_step": ns, "scan_size": 10, "stock_probes_per_step": n},
            "order_line_scan": {"rows_per_s": rows / (ms_scan * 1e-3), "scans_per_s": ns / (ms_scan * 1e-3), "ms_per_step": ms_scan,
                                "plan_ms": ol_plan_ms, "gather_ms": ol_gather_ms,
                                "achieved_GBps": rows * ol_row / (ms_scan * 1e-3) / 1e9, "frac_of_peak": rows * ol_row / (ms_scan * 1e-3) / 1e9 / peak},
            "stock_probe_full_row": {"probes_per_s": n / (ms_full * 1e-3), "ms_per_step": ms_full,
                                     "achieved_GBps": n * st_row / (ms_full * 1e-3) / 1e9, "frac_of_peak": n * st_row / (ms_full * 1e-3) / 1e9 / peak},
            "stock_probe_quantity_pushdown": {"probes_per_s": n / (ms_col * 1e-3), "ms_per_step": ms_col,
                                              "bytes_returned_per_probe": 5},
            "stock_probe_quantity_aggregate": {"probes_per_s": n / (ms_agg * 1e-3), "ms_per_step": ms_agg,
                                               "bytes_returned_per_batch": 24, "found": int(agg[0].found),
                                               "below_threshold_15": int(agg[0].below), "min_quantity": int(agg[0].min_value),
                                               "note": "count(S_QUANTITY < 15) and MIN(S_QUANTITY) reduced on the device; "
                                                       "includes the 24-byte D2H read and its synchronisation per batch"},
            "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s", "peak_source": peak_src}}
    print(json.dumps(line), flush=True)
    stock.close()
    ol.close()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.workload != "scan":
        run_extra_workload(args)
        return

    import torch
    import stage_indexorganized_b200 as sg

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    numa = None
    if world > 1 and not os.environ.get("STAGE_NO_NUMA_BIND"):
        from stage_indexorganized_b200.numa import bind_to_gpu_node
        numa = bind_to_gpu_node(local_rank)      # pinned e2e buffers local to this GPU's PCIe root
    dev = torch.device("cuda", local_rank)
    sg.init(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=dev)

    if world > 1:
        args.rows = args.rows_multi
    rows, S, n_scans, n_probes = args.rows, args.scan_size, args.scans, args.probes
    updates = rows // 10
    use_host_tree = world == 1 and args.table == "host"
    host_table, table_kind = None, None
    t_build0 = time.time()
    read_ts = {"first": FIRST_CID, "mid": FIRST_CID + updates // 2, "latest": FIRST_CID + updates}[args.read_ts]
    vis = 0 if args.table_scan else 1
    if not use_host_tree:
        # product-only construction: the table is generated on the device in the reference's row function and key
        # order, then the committed updates are installed by the refresh kernels (version chains on 10 % of rows)
        from stage_indexorganized_b200.ycsb import rank_to_key
        halo = 2 * S + 128
        total = rows * world
        first = rank * rows
        n_local = min(rows + (halo if rank + 1 < world else 0), total - first)
        snap = sg.Snapshot.generate_ycsb(total, P_BYTES, rows_per_leaf=61, first_rank=first, n_local=n_local)
        if world == 1:
            uk = update_keys(rows, updates)
        else:       # updates of keys this shard owns
            uk = rank_to_key(total, first + np.random.default_rng(77 + rank).integers(0, rows, updates))
        applied = apply_updates_on_device(snap, uk, torch, dev)
        assert applied == updates, (applied, updates)
        table_desc = ("generated on device in reference key order (consolidated leaves, 61 rows/leaf), "
                      f"{updates} committed updates installed by stage_gpu_snapshot_apply_updates")
    else:
        host_table, table_kind = build_host_table(rows, updates)
        snap = sg.Snapshot.from_tree(host_table.root, K_BYTES, P_BYTES)
        table_desc = f"host tree built by the reference loader ({table_kind}), flattened by the transform kernel"
    t_build = time.time() - t_build0
    info = snap.info()
    stream = torch.cuda.current_stream()
    snap.set_stream(stream.cuda_stream)

    # ---- device-resident inputs / outputs -------------------------------------------------------
    n_batches = args.warmup + args.steps
    if world > 1:
        rng = np.random.default_rng(1000 + rank)
        sk_all = rank_to_key(rows * world, rank * rows + rng.integers(0, rows, size=(n_batches, n_scans)))
        pk_all = rng.integers(0, rows * world + 1, size=(n_batches, n_probes)).astype(np.uint32)
    else:
        from stage_indexorganized_b200 import ycsb      # the drivers' generators (FastRandom), restated in the package
        sk_all = np.stack([scan_start_keys(rows, n_scans, seed=100 + b) for b in range(n_batches)])
        pk_all = np.stack([ycsb.uniform_keys(rows, n_probes, seed=200 + b) for b in range(n_batches)])
    d_sk = torch.from_numpy(sk_all.astype(np.uint32).view(np.int32)).to(dev)
    d_sizes = torch.full((n_scans,), S, dtype=torch.int32, device=dev)
    d_counts = torch.zeros((n_batches, n_scans), dtype=torch.int32, device=dev)
    d_keyout = torch.empty((n_scans * S,), dtype=torch.int32, device=dev)
    d_tsout = torch.empty((n_scans * S,), dtype=torch.int32, device=dev)
    d_payout = torch.empty((n_scans * S * P_BYTES,), dtype=torch.uint8, device=dev)

    def scan_step(b):
        snap.scan_device(d_sk[b], d_sizes, n_scans, S, read_ts, vis, d_counts[b], d_keyout, d_tsout, d_payout)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for b in range(args.warmup):
        scan_step(b)
    barrier()
    launches0 = sg.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with ClockSampler(local_rank) as clk:
        ev[0].record(stream)
        for i in range(args.steps):
            scan_step(args.warmup + i)
            ev[i + 1].record(stream)
        barrier()
    launches = sg.launch_count() - launches0
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    total_ms = ev[0].elapsed_time(ev[-1])
    rows_emitted = int(d_counts[args.warmup:].sum().item())
    # per-kernel durations of the same K steps (CUDA events recorded by the library around each launch)
    snap.enable_timing(True)
    plan_ms, gather_ms = [], []
    for i in range(args.steps):
        scan_step(args.warmup + i)
        p_, g_ = snap.last_timing()
        plan_ms.append(p_)
        gather_ms.append(g_)
    snap.enable_timing(False)

    # ---- probe leg (device resident) ------------------------------------------------------------
    probe = None
    if world == 1:
        d_pk = torch.from_numpy(pk_all.view(np.int32)).to(dev)
        d_found = torch.zeros((n_batches, n_probes), dtype=torch.uint8, device=dev)
        d_pts = torch.empty((n_probes,), dtype=torch.int32, device=dev)
        d_pkey = torch.empty((n_probes,), dtype=torch.int32, device=dev)
        d_ppay = d_payout[: n_probes * P_BYTES] if n_probes <= n_scans * S else torch.empty(
            (n_probes * P_BYTES,), dtype=torch.uint8, device=dev)
        for b in range(args.warmup):
            snap.probe_device(d_pk[b], n_probes, read_ts, d_found[b], d_pts, d_pkey, d_ppay)
        torch.cuda.synchronize()
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pe0.record(stream)
        for i in range(args.steps):
            b = args.warmup + i
            snap.probe_device(d_pk[b], n_probes, read_ts, d_found[b], d_pts, d_pkey, d_ppay)
        pe1.record(stream)
        torch.cuda.synchronize()
        pms = pe0.elapsed_time(pe1)
        snap.enable_timing(True)
        pp, pg = [], []
        for i in range(args.steps):
            b = args.warmup + i
            snap.probe_device(d_pk[b], n_probes, read_ts, d_found[b], d_pts, d_pkey, d_ppay)
            p_, g_ = snap.last_timing()
            pp.append(p_)
            pg.append(g_)
        snap.enable_timing(False)
        hits = int(d_found[args.warmup:].sum().item())
        nprobe = args.steps * n_probes
        peak, peak_src = peaks()
        pbytes = hits * PROBE_HIT_BYTES + (nprobe - hits) * PROBE_MISS_BYTES
        fused = sum(pp) < 0.02 * sum(pg)          # one warp-specialised kernel does search + rows (probe_fused.cuh)
        g_ach = (pbytes if fused else hits * 2 * P_BYTES) / (sum(pg) * 1e-3) / 1e9
        probe = {"value": nprobe / (pms * 1e-3), "unit": "probes/s", "ms_per_step": pms / args.steps,
                 "hit_rate": hits / nprobe, "plan_ms": sum(pp) / len(pp), "gather_ms": sum(pg) / len(pg),
                 "roofline": {"bound": "hbm", "kernel": "stage::probe_fused_kernel (search + MVCC + payload rows)" if fused
                              else "stage::gather_rows_kernel", "achieved": g_ach, "peak": peak,
                              "unit": "GB/s", "frac": g_ach / peak, "traffic": None,
                              "step_achieved": pbytes / (pms * 1e-3) / 1e9,
                              "step_frac": pbytes / (pms * 1e-3) / 1e9 / peak}}

    # ---- probe leg at N > 1: range-routed probes, exchange fused into the kernels (peer memory over NVLink) ----
    if world > 1:
        from stage_indexorganized_b200 import sharded
        split = sharded.ycsb_splitters(rows * world, world)
        d_pk = torch.from_numpy(pk_all.astype(np.uint32).view(np.uint8).reshape(n_batches, n_probes, K_BYTES)).to(dev)
        pp = sharded.PeerProbe(snap, split, n_probes, device_sync=True)

        def timed(fn, steps):
            for b in range(args.warmup):
                fn(b)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            hits_ = []
            for i in range(steps):
                f = fn(args.warmup + i)[0]
                hits_.append(f.sum())          # device-side reduction of this step's found flags (slab is reused)
            e1.record(stream)
            barrier()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            h = torch.stack(hits_).sum().to(torch.float64).reshape(1)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(h, op=dist.ReduceOp.SUM)
            return float(t.item()), float(h.item())

        l0 = sg.launch_count()
        pms, hit_sum = timed(lambda b: pp.probe(d_pk[b], n_probes, read_ts), args.steps)
        probe_launches = sg.launch_count() - l0
        if pp.x.error() != 0:
            raise SystemExit(f"probe exchange error {pp.x.error()} on rank {rank}")
        nprobe = args.steps * n_probes * world
        hit_rate = hit_sum / nprobe
        remote = (world - 1) / world
        # per GPU per step: HBM = local algorithmic bytes; NVLink (out direction) = rows served to remote requesters
        hbm_bytes = n_probes * (hit_rate * PROBE_HIT_BYTES + (1 - hit_rate) * PROBE_MISS_BYTES)
        nvl_bytes = n_probes * remote * (hit_rate * P_BYTES + 5 + K_BYTES + 4)
        peak, _ = peaks()
        nvl_peak = 770.0                                           # GB/s per direction per GPU, measured peer copy (B200_PROFILING.md)
        t_hbm, t_nvl = hbm_bytes / (peak * 1e9), nvl_bytes / (nvl_peak * 1e9)
        step_s = pms * 1e-3 / args.steps
        probe = {"value": nprobe / (pms * 1e-3), "unit": "probes/s", "ms_per_step": pms / args.steps,
                 "hit_rate": hit_rate, "launches_per_step": probe_launches // (args.steps + args.warmup),
                 "routing": "exchange fused into the kernels (csrc/exchange.cuh): keys stored into the owners' inboxes, "
                            "payload rows stored into the requesters' result slabs over NVLink peer memory (CUDA IPC), "
                            "device-side epoch flags, results in request order",
                 "nvlink_bytes_per_probe": nvl_bytes / n_probes,
                 "roofline": {"bound": "nvlink" if t_nvl > t_hbm else "hbm",
                              "target_ms": 1e3 * max(t_hbm, t_nvl), "frac": max(t_hbm, t_nvl) / step_s,
                              "nvlink_achieved": nvl_bytes / step_s / 1e9, "nvlink_peak": nvl_peak,
                              "hbm_achieved": hbm_bytes / step_s / 1e9, "hbm_peak": peak, "unit": "GB/s"}}
        pp.x.enable_timing(True)
        phases = []
        for i in range(args.steps):
            pp.probe(d_pk[args.warmup + i], n_probes, read_ts)
            phases.append(pp.x.last_timing())
        pp.x.enable_timing(False)
        probe["phase_ms_rank0"] = dict(zip(("post", "search_incl_wait_for_posts", "rows", "wait_for_owners"),
                                           [float(np.mean(c)) for c in zip(*phases)]))
        # comparison: the same probes with NCCL all-to-all collectives around the local kernels
        d_owner = torch.empty((n_probes,), dtype=torch.int32, device=dev)

        def local_probe(kin):
            m = kin.shape[0]
            f = torch.zeros((m,), dtype=torch.uint8, device=dev)
            t = torch.zeros((m,), dtype=torch.int32, device=dev)
            pay = torch.empty((m, P_BYTES), dtype=torch.uint8, device=dev)
            if m:
                snap.probe_device(kin, m, read_ts, f, t, None, pay)
            return f, t, pay

        def a2a_step(b):
            sg.route_device(d_pk[b], n_probes, split, d_owner, K_BYTES, stream.cuda_stream)
            return sharded.routed_probe(d_pk[b], d_owner, local_probe, P_BYTES)

        if args.a2a_steps > 0:
            ams, _ = timed(a2a_step, args.a2a_steps)
            probe["nccl_all_to_all"] = {"value": args.a2a_steps * n_probes * world / (ams * 1e-3), "unit": "probes/s",
                                        "ms_per_step": ams / args.a2a_steps}
        pp.close()

        # ---- routed scans: start keys anywhere in the global table, each scan planned on the owner of its start key
        # (shard + halo), rows stored into the requester's slab over NVLink peer memory -------------------------
        n_rs = args.routed_scans
        if n_rs > 0:
            rng2 = np.random.default_rng(3000 + rank)
            rk_all = rank_to_key(rows * world, rng2.integers(0, rows * world, size=(n_batches, n_rs)))
            d_rk = torch.from_numpy(rk_all.astype(np.uint32).view(np.uint8).reshape(n_batches, n_rs, K_BYTES)).to(dev)
            d_rsz = torch.full((n_rs,), S, dtype=torch.int32, device=dev)
            ps = sharded.PeerScan(snap, split, n_rs, S, device_sync=True)

            def rscan(b):
                c = ps.scan(d_rk[b], d_rsz, n_rs, read_ts)[0]
                return (c,)

            rms, rrows = timed(rscan, args.steps)
            if ps.x.error() != 0:
                raise SystemExit(f"scan exchange error {ps.x.error()} on rank {rank}")
            ps.x.enable_timing(True)
            phases = []
            for i in range(args.steps):
                ps.scan(d_rk[args.warmup + i], d_rsz, n_rs, read_ts)
                phases.append(ps.x.last_timing())
            ps.x.enable_timing(False)
            rstep = rms * 1e-3 / args.steps
            rows_step_gpu = rrows / args.steps / world
            routed_scans = {"value": rrows / (rms * 1e-3), "unit": "rows/s", "ms_per_step": rms / args.steps,
                            "scans_per_gpu_per_step": n_rs, "scan_size": S,
                            "nvlink_achieved": rows_step_gpu * remote * (P_BYTES + K_BYTES + 4) / rstep / 1e9,
                            "nvlink_peak": nvl_peak, "hbm_achieved": rows_step_gpu * ROW_ALGO_BYTES / rstep / 1e9,
                            "phase_ms_rank0": dict(zip(("post", "plan_incl_wait_for_posts", "rows", "wait_for_owners"),
                                                       [float(np.mean(c)) for c in zip(*phases)]))}
            probe["routed_scans"] = routed_scans
            ps.close()

    # ---- max over ranks, whole-job aggregate ----------------------------------------------------
    t_ms = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    r_sum = torch.tensor([rows_emitted], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(r_sum, op=dist.ReduceOp.SUM)
    total_ms_max = float(t_ms.item())
    rows_all = float(r_sum.item())
    value = rows_all / (total_ms_max * 1e-3)

    # ---- e2e: host buffers through the C ABI ----------------------------------------------------
    ne = min(args.e2e_scans, n_scans)
    h_sk = sg.PinnedArray((n_batches, ne, K_BYTES), np.uint8)
    h_sk.array[:] = sk_all[:, :ne].astype(np.uint32).view(np.uint8).reshape(n_batches, ne, K_BYTES)
    h_sizes = np.full(ne, S, np.uint32)
    h_counts = sg.PinnedArray((ne,), np.uint32)
    h_key = sg.PinnedArray((ne * S, K_BYTES), np.uint8)
    h_ts = sg.PinnedArray((ne * S,), np.uint32)
    h_pay = sg.PinnedArray((ne * S, P_BYTES), np.uint8)

    def e2e_step(b):
        sg._check(sg.lib().stage_gpu_scan_batch(snap.h, h_sk.array[b].ctypes.data, h_sizes.ctypes.data, ne, S, read_ts,
                                                vis, h_counts.ptr, h_key.ptr, h_ts.ptr, h_pay.ptr, sg.IO_HOST))
        return int(h_counts.array.sum())

    for b in range(min(args.warmup, 2)):
        e2e_step(b)
    barrier()
    t0 = time.perf_counter()
    e_rows = 0
    for i in range(args.steps):
        e_rows += e2e_step(args.warmup + i)
    torch.cuda.synchronize()
    e_sec = time.perf_counter() - t0
    e_t = torch.tensor([e_sec], dtype=torch.float64, device=dev)
    e_r = torch.tensor([e_rows], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e_r, op=dist.ReduceOp.SUM)
    e2e = {"value": float(e_r.item()) / float(e_t.item()), "unit": "rows/s",
           "h2d_bytes_per_step": ne * (K_BYTES + 4),
           "d2h_bytes_per_step": ne * 4 + ne * S * (K_BYTES + 4 + P_BYTES),
           "scans_per_step": ne, "rank0_seconds": e_sec, "numa_bind_rank0": numa}

    # ---- CPU baseline (rank 0, N = 1): the only part of the run that touches oracle/ ---------------------------
    cpu, transform, parity = None, None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        if host_table is None:
            host_table, table_kind = build_host_table(rows, updates)     # reference loader + the same updates
        threads = os.cpu_count() or 1
        n_cpu = args.cpu_scans or calibrated_scan_count(host_table, rows, S, read_ts, threads, 15.0)
        sk = scan_start_keys(rows, n_cpu, seed=100 + args.warmup).view(np.uint8).reshape(-1, 4)
        rps, r, sec = cpu_scan_sample(host_table, sk, S, read_ts, threads, vis)
        cpu = {"value": rps, "unit": "rows/s", "cores": threads, "kind": table_kind,
               "sample": f"{n_cpu} scans x {S} rows ({r} rows, {sec:.1f} s), {threads} pinned threads, "
                         "table built by the reference loader with the same rows and the same committed updates"}
        # with the reference-format tree at hand: time the record-layout transform on it and cross-check a sample
        t0 = time.time()
        hsnap = snap if use_host_tree else sg.Snapshot.from_tree(host_table.root, K_BYTES, P_BYTES)
        t_tr = time.time() - t0
        hi = hsnap.info()
        if not use_host_tree:
            transform = {"leaves": int(hi.n_leaves), "rows": int(hi.n_rows), "seconds": t_tr,
                         "leaf_image_GB_per_s": hi.n_leaves * 65536 / t_tr / 1e9,
                         "note": "stage_gpu_snapshot_from_tree: host walk + pinned staging + H2D + leaf_transform_kernel"}
        ns = 256
        got = hsnap.scan(sk[:ns], S, read_ts, vis)
        want = host_table.scan(sk[:ns], S, read_ts, vis)
        pk = pk_all[0][:65536].astype(np.uint32).view(np.uint8).reshape(-1, 4)
        gp, wp = snap.probe(pk, read_ts), host_table.probe(pk, read_ts)
        ok = all(np.array_equal(g, w) for g, w in zip(got, want)) and all(np.array_equal(g, w) for g, w in zip(gp, wp))
        parity = {"ok": bool(ok), "checked": f"{ns} scans x {S} rows (flattened host tree vs CPU {table_kind}) and "
                                             f"{len(pk)} probes (bench table vs CPU {table_kind}), bit-exact"}
        if not use_host_tree:
            # structural refresh: 10 k more committed updates on the host tree, then only the changed leaves are staged
            uk2 = np.random.default_rng(99).integers(0, rows, 10_000).astype(np.uint32).view(np.uint8).reshape(-1, 4)
            host_table.commit_update_batch(uk2, FIRST_CID + updates + 1000)
            t0 = time.time()
            nsnap, st = hsnap.refresh_from_tree(host_table.root)
            t_rf = time.time() - t0
            ts2 = FIRST_CID + updates + 6000
            ok2 = all(np.array_equal(g, w) for g, w in zip(nsnap.scan(sk[:64], S, ts2, 1), host_table.scan(sk[:64], S, ts2, 1)))
            transform["refresh_after_10k_updates"] = {
                "seconds": t_rf, "leaves_transformed": int(st.leaves_transformed), "leaves_carried": int(st.leaves_carried),
                "host_MB_staged": st.host_bytes_staged / 1e6, "parity_ok": bool(ok2),
                "note": "stage_gpu_snapshot_refresh_from_tree: unchanged leaves carried over on the device"}
            parity["ok"] = bool(parity["ok"] and ok2)
            nsnap.close()
        if hsnap is not snap:
            hsnap.close()

    if rank == 0:
        peak, peak_src = peaks()
        algo_bytes = rows_emitted * ROW_ALGO_BYTES + args.steps * n_scans * (K_BYTES + 8)
        kern_ms = sum(step_ms)
        step_achieved = algo_bytes / (kern_ms * 1e-3) / 1e9
        gather_bytes = rows_emitted * 2 * P_BYTES            # the gather kernel moves the payload rows only
        achieved = gather_bytes / (sum(gather_ms) * 1e-3) / 1e9
        line = {
            "metric": "YCSB scan rows/s (1000-row MVCC range scans, 10M-tuple int table)",
            "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms_max / args.steps, "step_ms_median": float(np.median(step_ms)),
            "step_ms_best": float(min(step_ms)), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args, world, table_desc),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_PER_LAUNCH, "peak_source": peak_src,
                         "kernel": "stage::gather_rows_kernel<uint2,4> (payload rows; dominant kernel of the step)",
                         "algorithmic_bytes_per_launch": gather_bytes / args.steps,
                         "launch_ms": sum(gather_ms) / args.steps,
                         "plan_kernel_ms": sum(plan_ms) / args.steps,
                         "step_algorithmic_bytes": algo_bytes / args.steps,
                         "step_achieved": step_achieved, "step_frac": step_achieved / peak,
                         "nominal_peak": 8000.0, "frac_of_nominal": achieved / 8000.0,
                         "step_frac_of_nominal": step_achieved / 8000.0},
            "cpu_baseline": cpu,
            "parity_check": parity,
            "transform": transform,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clk.summary(),
            "probe": probe,
            "snapshot": {"rows": int(info.n_rows), "leaves": int(info.n_leaves), "versions": int(info.n_versions),
                         "hbm_bytes": int(info.device_bytes), "build_s": t_build},
        }
        print(json.dumps(line), flush=True)
    snap.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
