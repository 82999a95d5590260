#!/usr/bin/env python
"""One short pass over the hot kernels for `ncu --set full` (profiles/README.md, round 2).

    python tools/profile_kernels.py                       # plain run (must exit 0 before the ncu run)
    ncu --set full --clock-control none --import-source on --profile-from-start off \
        -o gpurun_out/r2_prof python tools/profile_kernels.py

Every kernel of interest runs twice untimed (warm-up), then once between cudaProfilerStart / cudaProfilerStop:
  headline   scan_plan_kernel<1> + gather_rows_kernel<uint2,4>   4096 x 1000-row MVCC scans, 10 M-row YCSB table
  probes     probe_kernel<1,4,true> + gather_rows_kernel          2^20 uniform probes, same table
  config 3   scan_short_fused_kernel<8,16,*>                      10^6 order_line scans of 10 rows (TPC-C w = 30)
             probe_kernel<4,4,true> + gather_rows_mid_kernel<4>   2^20 stock probes, 400-byte rows
             stage_gpu_tpcc_stock_level                           50 000 stock-level transactions
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stage_indexorganized_b200 as sg                      # noqa: E402
from stage_indexorganized_b200 import tpcc, ycsb          # noqa: E402
import bench_workloads as bw                                # noqa: E402


def main():
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    sg.init(0)
    st = torch.cuda.current_stream().cuda_stream
    rows, S, n_scans, n_probes = 10_000_000, 1000, 4096, 1 << 20
    snap = sg.Snapshot.generate_ycsb(rows, 1000, rows_per_leaf=61)
    uk = (ycsb.uniform_keys(rows, rows // 10, seed=2) - 1).astype(np.uint32)
    d_uk = torch.from_numpy(uk.view(np.uint8).reshape(-1, 4)).to(dev)
    cids = torch.arange(10, 10 + len(uk), dtype=torch.int32, device=dev)
    patches = (cids & 0xFF).to(torch.uint8)[:, None].expand(len(uk), 100).contiguous()
    snap.apply_updates(d_uk, patches, cids, patch_off=100, device=True)
    snap.set_stream(st)
    read_ts = 10 + len(uk) // 2
    sk = torch.from_numpy((ycsb.uniform_keys(rows, n_scans, seed=100) - 1).astype(np.uint32).view(np.int32)).to(dev)
    sizes = torch.full((n_scans,), S, dtype=torch.int32, device=dev)
    counts = torch.zeros((n_scans,), dtype=torch.int32, device=dev)
    kout = torch.empty((n_scans * S,), dtype=torch.int32, device=dev)
    tso = torch.empty((n_scans * S,), dtype=torch.int32, device=dev)
    pout = torch.empty((n_scans * S * 1000,), dtype=torch.uint8, device=dev)
    pk = torch.from_numpy(ycsb.uniform_keys(rows, n_probes, seed=200).view(np.int32)).to(dev)
    found = torch.zeros((n_probes,), dtype=torch.uint8, device=dev)
    pts = torch.empty((n_probes,), dtype=torch.int32, device=dev)

    def ycsb_pass():
        snap.scan_device(sk, sizes, n_scans, S, read_ts, 1, counts, kout, tso, pout)
        snap.probe_device(pk, n_probes, read_ts, found, pts, kout, pout)

    W = 30
    skeys, spay = tpcc.stock_rows(W)
    okeys, opay, _ = tpcc.order_line_rows(W)
    stock = bw._bulk(sg, torch, skeys, spay, tpcc.STOCK_KEY, tpcc.STOCK_PAYLOAD)
    ol = bw._bulk(sg, torch, okeys, opay, tpcc.OL_KEY, tpcc.OL_PAYLOAD)
    txns = 50_000
    ns = txns * 20
    starts = torch.from_numpy(tpcc.stock_level_requests(W, txns, seed=10)[0]).to(dev)
    osz = torch.full((ns,), 10, dtype=torch.int32, device=dev)
    ocnt = torch.zeros((ns,), dtype=torch.int32, device=dev)
    okout = torch.empty((ns * 10, tpcc.OL_KEY), dtype=torch.uint8, device=dev)
    otso = torch.empty((ns * 10,), dtype=torch.int32, device=dev)
    opout = torch.empty((ns * 10, tpcc.OL_PAYLOAD), dtype=torch.uint8, device=dev)
    rng = np.random.default_rng(3)
    spk = torch.from_numpy(tpcc.stock_key(rng.integers(1, W + 1, n_probes), rng.integers(1, tpcc.ITEMS + 1, n_probes))).to(dev)
    sfound = torch.zeros((n_probes,), dtype=torch.uint8, device=dev)
    spout = torch.empty((n_probes, tpcc.STOCK_PAYLOAD), dtype=torch.uint8, device=dev)
    req = np.zeros(txns, sg.STOCK_LEVEL_REQ)
    req["w_id"] = rng.integers(1, W + 1, txns)
    req["d_id"] = rng.integers(1, tpcc.DISTRICTS + 1, txns)
    req["next_o_id"] = tpcc.ORDERS + 1
    req["threshold"] = rng.integers(10, 21, txns)
    d_req = torch.from_numpy(req.view(np.uint8).reshape(txns, 32)).to(dev)
    d_cnt = torch.zeros((txns,), dtype=torch.int32, device=dev)

    def tpcc_pass():
        ol.scan_device(starts, osz, ns, 10, 1, 1, ocnt, okout, otso, opout)
        stock.probe_device(spk, n_probes, 1, sfound, None, None, spout)
        sg.tpcc_stock_level(ol, stock, d_req, 1, out=d_cnt, n=txns)

    for _ in range(2):
        ycsb_pass()
        tpcc_pass()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    ycsb_pass()
    tpcc_pass()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("profiled pass done:", int(counts.sum().item()), "scan rows,", int(found.sum().item()), "probe hits,",
          int(ocnt.sum().item()), "order_line rows,", int(sfound.sum().item()), "stock hits,", float(d_cnt.float().mean().item()),
          "mean distinct items")


if __name__ == "__main__":
    main()
