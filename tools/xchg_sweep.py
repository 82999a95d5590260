#!/usr/bin/env python
"""Routed-probe exchange sweep (run under torchrun, one rank per GPU): the table is built once, then the row phase is
timed under the runtime knobs of the exchange (tile order, CTAs per SM of the row mover, TMA bulk-store mover, batch
size). One JSON line per variant from rank 0. Results are checked against the row function on every variant.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29577 \
        tools/xchg_sweep.py [--rows-per-gpu 125000000] [--steps 8]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stage_indexorganized_b200 as sg                      # noqa: E402
from stage_indexorganized_b200 import sharded               # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-per-gpu", type=int, default=125_000_000)
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    sg.init(local)
    dist.init_process_group("nccl", device_id=dev)
    rows, P = args.rows_per_gpu, 1000
    total = rows * world
    snap = sg.Snapshot.generate_ycsb(total, P, rows_per_leaf=61, first_rank=rank * rows, n_local=rows)
    stream = torch.cuda.current_stream()
    snap.set_stream(stream.cuda_stream)
    split = sharded.ycsb_splitters(total, world)
    rng = np.random.default_rng(1000 + rank)
    variants = [("baseline", {}, 1 << 20)]
    for cps in (1, 2):
        variants.append((f"ctas_per_sm={cps}", {"STAGE_XCHG_CTAS_PER_SM": str(cps)}, 1 << 20))
    variants.append(("order=permuted", {"STAGE_XCHG_ORDER": "0"}, 1 << 20))
    variants.append(("tma", {"STAGE_XCHG_TMA": "1"}, 1 << 20))
    variants.append(("tma ctas_per_sm=1", {"STAGE_XCHG_TMA": "1", "STAGE_XCHG_CTAS_PER_SM": "1"}, 1 << 20))
    variants.append(("tma order=permuted", {"STAGE_XCHG_TMA": "1", "STAGE_XCHG_ORDER": "0"}, 1 << 20))
    variants.append(("batch=2^22", {}, 1 << 22))
    variants.append(("tma batch=2^22", {"STAGE_XCHG_TMA": "1"}, 1 << 22))
    knobs = ("STAGE_XCHG_CTAS_PER_SM", "STAGE_XCHG_ORDER", "STAGE_XCHG_TMA")
    pp0 = sharded.PeerProbe(snap, split, 1 << 20, device_sync=True)
    link = pp0.measure_link_ceiling()
    pp0.close()
    if rank == 0:
        print(json.dumps({"variant": "link ceiling: peer copies (copy engines), all ranks at once, GB/s per direction per GPU",
                          "world": world, "GBps": link}), flush=True)
    for name, env, n in variants:
        for k in knobs:
            os.environ.pop(k, None)
        os.environ.update(env)
        keys_np = rng.integers(0, total + 1, size=(4, n)).astype(np.uint32)
        d_keys = torch.from_numpy(keys_np.view(np.uint8).reshape(4, n, 4)).to(dev)
        pp = sharded.PeerProbe(snap, split, n, device_sync=True)
        for b in range(3):
            pp.probe(d_keys[b], n, 0)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(args.steps):
            found, ts, pay = pp.probe(d_keys[i % 4], n, 0)
        e1.record(stream)
        torch.cuda.synchronize()
        dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        pp.x.enable_timing(True)
        ph = []
        for i in range(4):
            pp.probe(d_keys[i % 4], n, 0)
            ph.append(pp.x.last_timing())
        pp.x.enable_timing(False)
        # correctness of the last batch (keys d_keys[3]) against the row function
        kk = keys_np[3]
        f = found.cpu().numpy()
        want = (kk < total).astype(np.uint8)
        p0, p999 = pay[:, 0].cpu().numpy(), pay[:, 999].cpu().numpy()
        hit = want == 1
        ok = bool(np.array_equal(f, want) and np.array_equal(p0[hit], (kk[hit] & 0xFF).astype(np.uint8)) and
                  np.array_equal(p999[hit], (kk[hit] & 0xFF).astype(np.uint8)) and pp.x.error() == 0)
        okt = torch.tensor([int(ok)], dtype=torch.int32, device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        if rank == 0:
            phm = [float(np.mean(c)) for c in zip(*ph)]
            remote = (world - 1) / world
            print(json.dumps({"variant": name, "world": world, "probes_per_gpu": n, "ms_per_step": float(ms.item()),
                              "probes_per_s": n * world / (float(ms.item()) * 1e-3), "ok": bool(okt.item()),
                              "phase_ms_rank0": dict(zip(("post", "search", "rows", "wait"), phm)),
                              "row_phase_nvlink_GBps": n * remote * 1000 / (phm[2] * 1e-3) / 1e9 if world > 1 else None}),
                  flush=True)
        pp.close()
    snap.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
