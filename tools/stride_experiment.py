#!/usr/bin/env python
"""Row-stride experiment: does padding the 1000-byte YCSB rows to a 16-byte (1008) or
128-byte (1024) multiple pay for the payload gather?

The padded stride is emulated with a table whose payload_size IS the padded size (generate_ycsb takes any size): the
gather then moves rows of that size with the vector width the alignment allows (1000: 8-byte accesses, rows straddle
32-byte sectors; 1008: 16-byte accesses; 1024: 16-byte accesses, sector- and line-aligned rows). The useful bytes are
counted as for 1000-byte rows in every case, so a padded variant only wins if the better access shape more than pays for
the extra bytes it moves. Output: one JSON line per variant (HBM footprint, probe and scan step times, useful GB/s and
fraction of the measured copy peak).
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stage_indexorganized_b200 as sg          # noqa: E402
from stage_indexorganized_b200 import ycsb     # noqa: E402
import bench                                    # noqa: E402


def timed(fn, warm=3, steps=10):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    sg.init(0)
    peak, _ = bench.peaks()
    rows, S, n_scans, n_probes, nb = 10_000_000, 1000, 4096, 1 << 20, 13
    uk = (ycsb.uniform_keys(rows, rows // 10, seed=2) - 1).astype(np.uint32)
    sk = torch.from_numpy(np.stack([(ycsb.uniform_keys(rows, n_scans, seed=100 + b) - 1) for b in range(nb)]).astype(np.uint32).view(np.int32)).to(dev)
    pk = torch.from_numpy(np.stack([ycsb.uniform_keys(rows, n_probes, seed=200 + b) for b in range(nb)]).view(np.int32)).to(dev)
    for P in (1000, 1008, 1024):
        snap = sg.Snapshot.generate_ycsb(rows, P, rows_per_leaf=61)
        d_uk = torch.from_numpy(uk.view(np.uint8).reshape(-1, 4)).to(dev)
        cids = torch.arange(10, 10 + len(uk), dtype=torch.int32, device=dev)
        patches = (cids & 0xFF).to(torch.uint8)[:, None].expand(len(uk), 100).contiguous()
        snap.apply_updates(d_uk, patches, cids, patch_off=100, device=True)
        snap.set_stream(torch.cuda.current_stream().cuda_stream)
        read_ts = 10 + len(uk) // 2
        sizes = torch.full((n_scans,), S, dtype=torch.int32, device=dev)
        counts = torch.zeros((nb, n_scans), dtype=torch.int32, device=dev)
        kout = torch.empty((n_scans * S,), dtype=torch.int32, device=dev)
        tso = torch.empty((n_scans * S,), dtype=torch.int32, device=dev)
        pout = torch.empty((n_scans * S * P,), dtype=torch.uint8, device=dev)
        found = torch.zeros((nb, n_probes), dtype=torch.uint8, device=dev)
        pts = torch.empty((n_probes,), dtype=torch.int32, device=dev)
        ms_scan = timed(lambda b: snap.scan_device(sk[b], sizes, n_scans, S, read_ts, 1, counts[b], kout, tso, pout))
        ms_probe = timed(lambda b: snap.probe_device(pk[b], n_probes, read_ts, found[b], pts, kout, pout))
        n_rows_scan = float(counts[3:].sum().item()) / 10
        hits = float(found[3:].sum().item()) / 10
        useful_scan = n_rows_scan * 2 * (4 + 1000 + 4)
        useful_probe = hits * (4 + 2 * (4 + 1000 + 4))
        print(json.dumps({"payload_stride": P, "hbm_bytes": int(snap.info().device_bytes),
                          "scan_ms": ms_scan, "scan_useful_GBps": useful_scan / (ms_scan * 1e-3) / 1e9,
                          "scan_frac_of_measured": useful_scan / (ms_scan * 1e-3) / 1e9 / peak,
                          "probe_ms": ms_probe, "probe_useful_GBps": useful_probe / (ms_probe * 1e-3) / 1e9,
                          "probe_frac_of_measured": useful_probe / (ms_probe * 1e-3) / 1e9 / peak}), flush=True)
        snap.close()
        del pout
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
