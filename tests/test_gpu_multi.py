"""GPU, N >= 2 devices: one process per GPU, NCCL for the control plane, CUDA IPC peer memory for the probe exchange.
Skipped on single-GPU boxes (the same kernels are covered there by tests/test_gpu_exchange.py)."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_ROWS, P = 400_000, 1000


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import stage_indexorganized_b200 as sg
    from stage_indexorganized_b200 import sharded
    sg.init(rank)
    lo, hi = sharded.shard_bounds(N_ROWS, world)[rank]
    split = sharded.ycsb_splitters(N_ROWS, world)
    shard = sg.Snapshot.generate_ycsb(N_ROWS, P, rows_per_leaf=61, first_rank=lo, n_local=hi - lo)
    shard.set_stream(torch.cuda.current_stream().cuda_stream)
    ok = True
    rng = np.random.default_rng(50 + rank)
    n = 60_000 + 1000 * rank
    for device_sync in (False, True):
        pp = sharded.PeerProbe(shard, split, 70_000, device_sync=device_sync)
        for rep in range(3):
            keys_np = rng.integers(0, N_ROWS + 100, n).astype(np.uint32)
            keys = torch.from_numpy(keys_np.view(np.uint8).reshape(-1, 4).copy()).cuda()
            found, ts, pay = pp.probe(keys, n, 1)
            torch.cuda.synchronize()
            assert pp.x.error() == 0
            f = found.cpu().numpy()
            p = pay.cpu().numpy()
            want_f = (keys_np < N_ROWS).astype(np.uint8)          # row function of the YCSB loader
            ok = ok and np.array_equal(f, want_f)
            hit = want_f == 1
            ok = ok and np.array_equal(p[hit, 0], (keys_np[hit] & 0xFF).astype(np.uint8))
            ok = ok and np.array_equal(p[hit, P - 1], (keys_np[hit] & 0xFF).astype(np.uint8))
            ok = ok and bool((p[hit] == p[hit, :1]).all())
            # against the NCCL all-to-all form of the same probe
            owner = torch.empty((n,), dtype=torch.int32, device="cuda")
            sg.route_device(keys, n, split, owner, 4, torch.cuda.current_stream().cuda_stream)

            def local_probe(kin):
                m = kin.shape[0]
                fo = torch.zeros((m,), dtype=torch.uint8, device="cuda")
                t = torch.zeros((m,), dtype=torch.int32, device="cuda")
                py = torch.zeros((m, P), dtype=torch.uint8, device="cuda")
                if m:
                    shard.probe_device(kin, m, 1, fo, t, None, py)
                return fo, t, py

            f2, t2, p2 = sharded.routed_probe(keys, owner, local_probe, P)
            ok = ok and bool(torch.equal(found, f2)) and bool(torch.equal(pay[found.bool()], p2[found.bool()]))
            # routed push-down over real NVLink (last: it reuses the result slab that `pay` is a view of): 4 bytes per probe back, then one 24-byte record per (owner, requester)
            f3, _, c3 = pp.probe_column(keys, n, 1, P - 4, 4)
            torch.cuda.synchronize()
            c3 = c3.cpu().numpy()
            ok = ok and np.array_equal(f3.cpu().numpy(), want_f) and bool((c3[hit] == (keys_np[hit] & 0xFF).astype(np.uint8)[:, None]).all())
            ok = ok and not c3[~hit].any()
            agg = pp.probe_aggregate_i32(keys, n, 1, 0, 0x40404040)
            v = ((keys_np[hit] & 0xFF).astype(np.int64) * 0x01010101).astype(np.uint32).view(np.int32)
            ok = ok and (agg.found, agg.below, agg.min_value, agg.max_value) == (int(hit.sum()), int((v < 0x40404040).sum()),
                                                                                  int(v.min()), int(v.max()))
        pp.close()
    shard.close()
    ret[rank] = bool(ok)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_peer_probe_multi_process():
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 8)
    port = 29700 + (os.getpid() % 2000)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret.get(r, False) for r in range(world)), dict(ret)


def _scan_worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import stage_indexorganized_b200 as sg
    from stage_indexorganized_b200 import sharded
    sg.init(rank)
    RPL, S, n = 61, 200, 3000 + 10 * rank
    N = RPL * 8 * 800                                   # shard starts fall on leaf boundaries for world <= 8
    lo, hi = sharded.shard_bounds(N, world)[rank]
    split = sharded.ycsb_splitters(N, world)
    shard = sg.Snapshot.generate_ycsb(N, P, rows_per_leaf=RPL, first_rank=lo, n_local=min(hi + 2 * S + 128, N) - lo)
    full = sg.Snapshot.generate_ycsb(N, P, rows_per_leaf=RPL)   # the unsharded table, as the expected answer
    st = torch.cuda.current_stream().cuda_stream
    shard.set_stream(st)
    full.set_stream(st)
    ok = lo % RPL == 0
    rng = np.random.default_rng(90 + rank)
    for device_sync in (False, True):
        ps = sharded.PeerScan(shard, split, 4000, S, device_sync=device_sync)
        for rep in range(2):
            keys = torch.from_numpy(rng.integers(0, N + 100, n).astype(np.uint32).view(np.uint8).reshape(-1, 4).copy()).cuda()
            sizes = torch.from_numpy(rng.integers(0, S + 1, n).astype(np.int32)).cuda()
            cnt, ko, to, po = ps.scan(keys, sizes, n, 1)
            torch.cuda.synchronize()
            ok = ok and ps.x.error() == 0
            wc = torch.zeros(n, dtype=torch.int32, device="cuda")
            wk = torch.zeros((n, S, 4), dtype=torch.uint8, device="cuda")
            wt = torch.zeros((n, S), dtype=torch.int32, device="cuda")
            wp = torch.zeros((n, S, P), dtype=torch.uint8, device="cuda")
            full.scan_device(keys, sizes, n, S, 1, True, wc, wk, wt, wp)
            torch.cuda.synchronize()
            m = torch.arange(S, device="cuda")[None, :] < wc[:, None]
            ok = ok and bool(torch.equal(cnt, wc)) and bool(torch.equal(ko[m], wk[m]))
            ok = ok and bool(torch.equal(to[m], wt[m])) and bool(torch.equal(po[m], wp[m]))
        ps.close()
    shard.close()
    full.close()
    ret[rank] = bool(ok)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_peer_scan_multi_process():
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 8)
    port = 31700 + (os.getpid() % 2000)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_scan_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret.get(r, False) for r in range(world)), dict(ret)
