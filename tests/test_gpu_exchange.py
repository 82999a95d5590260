"""GPU: routed probes over peer memory (csrc/exchange.cuh) — post / serve kernels, request-order result slabs.

One process drives every rank of a small world on ONE device (arenas attached by pointer), so the whole routing
path — owner computation in reference key order, inbox slot claims, per-source segments, MVCC on the owner,
payload rows written into the requester's slab — is checked bit-exactly against the single-table CPU checker
without needing several GPUs. The multi-process / NVLink form is tests/test_gpu_multi.py."""
import numpy as np
import pytest
import torch

import stage_indexorganized_b200 as sg
from stage_indexorganized_b200 import sharded
from stage_indexorganized_b200.ycsb import rank_to_key
from helpers import u32_keys
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def _run_world(world, shards, split, key_batches, read_ts, K, P, flags):
    cap = max(len(k) for k in key_batches)
    xs = [sg.ProbeExchange(world, r, cap, K, P) for r in range(world)]
    try:
        arenas = [x.arena() for x in xs]
        for x in xs:
            x.attach_ptrs(arenas)
        dkeys = [torch.from_numpy(np.ascontiguousarray(k)).cuda() for k in key_batches]
        out = []
        for rep in range(2):                     # two batches: epochs, counter reset, slab reuse
            for r in range(world):
                xs[r].post(split, dkeys[r], len(key_batches[r]), flags)
            for r in range(world):
                xs[r].serve(shards[r], read_ts, flags)
            for r in range(world):
                if flags:
                    xs[r].wait()
            torch.cuda.synchronize()
            out = []
            for r in range(world):
                assert xs[r].error() == 0
                f, t, p = xs[r].results_torch(len(key_batches[r]))
                out.append((f.cpu().numpy().copy(), t.cpu().numpy().view(np.uint32).copy(), p.cpu().numpy().copy()))
        return out
    finally:
        for x in xs:
            x.close()


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("flags", [0, sg.XCHG_DEVICE_SYNC])
def test_routed_probe_generated_shards(world, flags):
    N, P = 40000, 1000
    sg.init(0)
    full = pyoracle.OracleTable(P)
    assert full.load_ycsb_int(0, N) == N
    bounds = sharded.shard_bounds(N, world)
    split = sharded.ycsb_splitters(N, world) if world > 1 else None
    shards = [sg.Snapshot.generate_ycsb(N, P, rows_per_leaf=61, first_rank=lo, n_local=hi - lo) for lo, hi in bounds]
    try:
        rng = np.random.default_rng(world)
        batches = [u32_keys(rng.integers(0, N + 40, 5000 + 17 * r)) for r in range(world)]
        got = _run_world(world, shards, split, batches, 1, 4, P, flags)
        for r in range(world):
            wf, wt, _, wp = full.probe(batches[r], 1)
            f, t, p = got[r]
            assert np.array_equal(f, wf), f"rank {r} found"
            assert np.array_equal(t[wf == 1], wt[wf == 1])
            assert np.array_equal(p[wf == 1], wp[wf == 1]), f"rank {r} payload"
    finally:
        for s in shards:
            s.close()


def test_routed_probe_mvcc_shards_from_trees():
    """Shards flattened from checker-built trees with committed updates (version chains on the owners)."""
    N, P, world = 6000, 64, 2
    sg.init(0)
    full = pyoracle.OracleTable(P)
    full.load_ycsb_int(0, N)
    rng = np.random.default_rng(3)
    upd = u32_keys(rng.integers(0, N, 1500))
    assert full.commit_update_batch(upd, 10) == 1500
    bounds = sharded.shard_bounds(N, world)
    split = sharded.ycsb_splitters(N, world)
    order = rank_to_key(N, np.arange(N)).astype(np.uint32)
    owner_of = {int(k): (0 if i < bounds[0][1] else 1) for i, k in enumerate(order)}
    tables, shards = [], []
    for lo, hi in bounds:
        t = pyoracle.OracleTable(P)
        mine = order[lo:hi]
        pay = np.empty((len(mine), P), np.uint8)
        pay[:] = (mine & 0xFF).astype(np.uint8)[:, None]
        t.insert(u32_keys(mine), pay, 0)
        tables.append(t)
    for i, k in enumerate(upd.view(np.uint32).ravel()):       # same updates, same commit ids, on the owning shard
        assert tables[owner_of[int(k)]].commit_update_batch(upd[i:i + 1], 10 + i) == 1
    shards = [sg.Snapshot.from_tree(t.root, 4, P) for t in tables]
    try:
        batches = [u32_keys(rng.integers(0, N + 10, 3000)) for _ in range(world)]
        for read_ts in (0, 9, 700, 1509, 4000):
            got = _run_world(world, shards, split, batches, read_ts, 4, P, sg.XCHG_DEVICE_SYNC)
            for r in range(world):
                wf, wt, _, wp = full.probe(batches[r], read_ts)
                f, t, p = got[r]
                assert np.array_equal(f, wf), f"ts={read_ts} rank {r} found"
                assert np.array_equal(t[wf == 1], wt[wf == 1]), f"ts={read_ts} stamps"
                assert np.array_equal(p[wf == 1], wp[wf == 1]), f"ts={read_ts} payload"
    finally:
        for s in shards:
            s.close()


def test_exchange_overflow_is_reported():
    sg.init(0)
    with sg.Snapshot.generate_ycsb(1000, 64) as shard:
        x = sg.ProbeExchange(1, 0, 100, 4, 64)
        try:
            keys = torch.zeros((200, 4), dtype=torch.uint8, device="cuda")
            with pytest.raises(sg.StageError):
                x.post(None, keys, 200)          # n > capacity is refused on the host
        finally:
            x.close()


# ---- routed scans ------------------------------------------------------------------------------------------
def _scan_full(full, keys, sizes, S, read_ts, vis, K, P):
    n = len(keys)
    dk = torch.from_numpy(np.ascontiguousarray(keys)).cuda()
    ds = torch.from_numpy(np.ascontiguousarray(sizes, np.uint32).view(np.int32)).cuda()
    cnt = torch.zeros(n, dtype=torch.int32, device="cuda")
    ko = torch.zeros((n, S, K), dtype=torch.uint8, device="cuda")
    to = torch.zeros((n, S), dtype=torch.int32, device="cuda")
    po = torch.zeros((n, S, P), dtype=torch.uint8, device="cuda")
    full.scan_device(dk, ds, n, S, read_ts, vis, cnt, ko, to, po)
    torch.cuda.synchronize()
    return cnt.cpu().numpy(), ko.cpu().numpy(), to.cpu().numpy(), po.cpu().numpy()


@pytest.mark.parametrize("world,S", [(1, 100), (2, 100), (3, 100), (8, 100), (2, 1000), (8, 10)])
@pytest.mark.parametrize("flags", [0, sg.XCHG_DEVICE_SYNC])
def test_routed_scans_match_unsharded_table(world, S, flags):
    """Scans posted to the owner of their start key, planned there (shard + halo of its right neighbour) and written
    into the requester's slab == the same scans on the unsharded table (same leaf boundaries: the shard starts are
    multiples of the leaf fill), with version chains on both sides."""
    RPL, P, K = 61, 200, 4
    N = RPL * 720
    halo = 2 * S + 128
    sg.init(0)
    bounds = sharded.shard_bounds(N, world)
    assert all(lo % RPL == 0 for lo, _ in bounds)
    split = sharded.ycsb_splitters(N, world) if world > 1 else None
    full = sg.Snapshot.generate_ycsb(N, P, rows_per_leaf=RPL)
    shards = [sg.Snapshot.generate_ycsb(N, P, rows_per_leaf=RPL, first_rank=lo, n_local=min(hi + halo, N) - lo)
              for lo, hi in bounds]
    xs = []
    try:
        rng = np.random.default_rng(100 * world + S)
        nu = 6000
        ukeys = rank_to_key(N, rng.integers(0, N, nu)).astype(np.uint32).view(np.uint8).reshape(-1, 4)
        patches = rng.integers(0, 256, (nu, 16), dtype=np.uint8)
        cids = np.arange(10, 10 + nu, dtype=np.uint32)
        assert full.apply_updates(ukeys, patches, cids, patch_off=8) == nu
        for s in shards:
            s.apply_updates(ukeys, patches, cids, patch_off=8)
        read_ts = 10 + nu // 2
        nscan = [300 + 7 * r for r in range(world)]
        batches = [u32_keys(rng.integers(0, N + 40, m)) for m in nscan]
        sizes = [rng.integers(0, S + 1, m).astype(np.uint32) for m in nscan]
        for sz in sizes:
            sz[:3] = (S, 1, 0)
        cap = max(nscan)
        xs = [sg.ProbeExchange(world, r, cap, K, P, max_scan_size=S) for r in range(world)]
        arenas = [x.arena() for x in xs]
        for x in xs:
            x.attach_ptrs(arenas)
        dk = [torch.from_numpy(np.ascontiguousarray(b)).cuda() for b in batches]
        dsz = [torch.from_numpy(z.view(np.int32)).cuda() for z in sizes]
        for vis in (True, False):
            for r in range(world):
                xs[r].post_scans(split, dk[r], dsz[r], nscan[r], flags)
            for r in range(world):
                xs[r].serve_scans(shards[r], read_ts, vis, flags)
            for r in range(world):
                if flags:
                    xs[r].wait()
            torch.cuda.synchronize()
            for r in range(world):
                assert xs[r].error() == 0
                cnt, ko, to, po = (t.cpu().numpy() for t in xs[r].scan_results_torch(nscan[r]))
                wc, wk, wt, wp = _scan_full(full, batches[r], sizes[r], S, read_ts, vis, K, P)
                assert np.array_equal(cnt, wc), f"rank {r} counts (vis={vis})"
                if vis:
                    assert (cnt < sizes[r]).any()          # the filter drops rows somewhere
                m = np.arange(S)[None, :] < wc[:, None]
                assert np.array_equal(ko[m], wk[m]), f"rank {r} keys"
                assert np.array_equal(to[m], wt[m]), f"rank {r} ts"
                assert np.array_equal(po[m], wp[m]), f"rank {r} payload"
    finally:
        for x in xs:
            x.close()
        for s in shards:
            s.close()
        full.close()


def test_scan_exchange_argument_errors():
    sg.init(0)
    with pytest.raises(sg.StageError):
        sg.ProbeExchange(1, 0, 16, 6, 64, max_scan_size=10)      # key size not a multiple of 4
    x = sg.ProbeExchange(1, 0, 16, 4, 64)
    s = sg.Snapshot.generate_ycsb(1000, 64)
    try:
        k = torch.zeros((4, 4), dtype=torch.uint8, device="cuda")
        z = torch.ones(4, dtype=torch.int32, device="cuda")
        with pytest.raises(sg.StageError):
            x.post_scans(None, k, z, 4)                          # probe exchange does not take scans
        with pytest.raises(sg.StageError):
            x.serve_scans(s, 1)
    finally:
        x.close()
        s.close()


def _run_world_pushdown(world, shards, split, key_batches, read_ts, K, P, flags, col_off, col_len, threshold):
    """post / serve_project / wait and post / serve_aggregate_i32 / wait for every rank of a one-device world.
    -> per rank: (found, ts, col[n, col_len]) and AggI32."""
    cap = max(len(k) for k in key_batches)
    xs = [sg.ProbeExchange(world, r, cap, K, P) for r in range(world)]
    try:
        arenas = [x.arena() for x in xs]
        for x in xs:
            x.attach_ptrs(arenas)
        dkeys = [torch.from_numpy(np.ascontiguousarray(k)).cuda() for k in key_batches]
        cols, aggs = [], []
        for rep in range(2):
            for r in range(world):
                xs[r].post(split, dkeys[r], len(key_batches[r]), flags)
            for r in range(world):
                xs[r].serve_project(shards[r], read_ts, col_off, col_len, flags)
            for r in range(world):
                if flags:
                    xs[r].wait()
            torch.cuda.synchronize()
            cols = []
            for r in range(world):
                assert xs[r].error() == 0
                f, t, c = xs[r].project_results_torch(len(key_batches[r]), col_len)
                cols.append((f.cpu().numpy().copy(), t.cpu().numpy().view(np.uint32).copy(), c.cpu().numpy().copy()))
            for r in range(world):
                xs[r].post(split, dkeys[r], len(key_batches[r]), flags)
            for r in range(world):
                xs[r].serve_aggregate_i32(shards[r], read_ts, col_off, threshold, flags)
            for r in range(world):
                if flags:
                    xs[r].wait()
            torch.cuda.synchronize()
            aggs = [xs[r].aggregate_result() for r in range(world)]
        return cols, aggs
    finally:
        for x in xs:
            x.close()


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("flags", [0, sg.XCHG_DEVICE_SYNC])
def test_routed_pushdown_matches_unsharded_table(world, flags):
    """Routed column projection and routed aggregate (SURVEY 8f f3 over the exchange): the column bytes and the
    (found, below, min, max) record every requester gets equal what the unsharded table gives for its batch — version
    rows served from the owners' chains included."""
    N, P = 9000, 64
    sg.init(0)
    full = pyoracle.OracleTable(P)
    full.load_ycsb_int(0, N)
    rng = np.random.default_rng(30 + world)
    upd = u32_keys(rng.integers(0, N, 2000))
    assert full.commit_update_batch(upd, 10) == 2000
    bounds = sharded.shard_bounds(N, world)
    split = sharded.ycsb_splitters(N, world) if world > 1 else None
    order = rank_to_key(N, np.arange(N)).astype(np.uint32)
    rank_of = np.empty(N, np.int64)
    rank_of[order.astype(np.int64)] = np.arange(N)
    tables = []
    for lo, hi in bounds:
        t = pyoracle.OracleTable(P)
        mine = order[lo:hi]
        pay = np.empty((len(mine), P), np.uint8)
        pay[:] = (mine & 0xFF).astype(np.uint8)[:, None]
        t.insert(u32_keys(mine), pay, 0)
        tables.append(t)
    ends = np.array([hi for _, hi in bounds])
    for i, k in enumerate(upd.view(np.uint32).ravel()):
        own = int(np.searchsorted(ends, rank_of[int(k)], "right"))
        assert tables[own].commit_update_batch(upd[i:i + 1], 10 + i) == 1
    shards = [sg.Snapshot.from_tree(t.root, 4, P) for t in tables]
    try:
        batches = [u32_keys(rng.integers(0, N + 30, 2500 + 11 * r)) for r in range(world)]
        for read_ts, col_off, col_len, thr in ((0, 0, 4, 0x30303030), (900, 3, 7, 0), (5000, 60, 4, 0x7F000000)):
            cols, aggs = _run_world_pushdown(world, shards, split, batches, read_ts, 4, P, flags, col_off, col_len, thr)
            for r in range(world):
                wf, wt, _, wp = full.probe(batches[r], read_ts)
                f, t, c = cols[r]
                assert np.array_equal(f, wf) and np.array_equal(t[wf == 1], wt[wf == 1]), f"rank {r} found / stamps"
                assert np.array_equal(c[wf == 1], wp[wf == 1][:, col_off:col_off + col_len]) and not c[wf == 0].any(), f"rank {r} column"
                v = np.ascontiguousarray(wp[wf == 1][:, col_off:col_off + 4]).view("<i4").ravel()
                a = aggs[r]
                assert (a.found, a.below) == (int(wf.sum()), int((v < thr).sum())), f"rank {r} aggregate counts"
                assert (a.min_value, a.max_value) == (int(v.min()), int(v.max())), f"rank {r} min / max"
    finally:
        for s in shards:
            s.close()


def test_tma_row_mover_variant():
    """STAGE_XCHG_TMA=1: payload rows leave through cp.async.bulk shared -> global stores (one bulk write per row) instead
    of per-lane 16-byte stores. Same results: the routed probe / scan tests of this file re-run in a child process."""
    import os
    import subprocess
    import sys
    if os.environ.get("STAGE_VARIANT_RUN"):
        pytest.skip("already inside a variant run")
    env = dict(os.environ, STAGE_XCHG_TMA="1", STAGE_VARIANT_RUN="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k",
                        "routed_probe or routed_scans"], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
