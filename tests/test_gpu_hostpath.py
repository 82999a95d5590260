"""Host-buffer (STAGE_IO_HOST) paths of the batch calls: results go through double-buffered HBM scratch + DMA copies
(default; strided or STAGE_IO_COMPACT layout) or — STAGE_HOST_PATH=zerocopy, re-run in a child process — are stored by
the kernels straight into pinned outputs. All must return what the CPU checker returns; concurrent callers share one
snapshot like the reference's pinned worker threads share one BTree
(/root/reference/benchmark/ycsb/ycsb_workload.cpp:522-529)."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from helpers import assert_probe_equal, assert_scan_equal, random_table, u32_keys, ycsb_int_table
from oracle import pyoracle

pytestmark = pytest.mark.gpu
ZEROCOPY = os.environ.get("STAGE_HOST_PATH") == "zerocopy"      # forced zero copy refuses pageable outputs


def _checker():
    return pyoracle.RefTable if pyoracle.have_ref() else pyoracle.OracleTable


@pytest.fixture(scope="module")
def mvcc_table():
    """20 k YCSB rows, 3 k committed updates (chains), a few deletes: leaves with unsorted suffixes from the loader."""
    t = ycsb_int_table(_checker(), 20000)
    uk = u32_keys(pyoracle.uniform_keys(20000, 3000, seed=2) - 1)
    assert t.commit_update_batch(uk, 10) == 3000
    for k in (17, 4242, 19999):
        t.delete(u32_keys(np.array([k], np.uint32))[0])
    snap = sg.Snapshot.from_tree(t.root, 4, 1000)
    yield t, snap
    snap.close()


def _pinned_scan(snap, sk, sizes, stride, read_ts, vis, want_payload=True):
    n, K, P = sk.shape[0], snap.schema.key_size, snap.schema.payload_size
    hk = sg.PinnedArray((max(n, 1), K), np.uint8)
    hk.array[:n] = sk
    sz = np.ascontiguousarray(np.broadcast_to(np.asarray(sizes, np.uint32), (n,)))
    counts = sg.PinnedArray((max(n, 1),), np.uint32)
    keys = sg.PinnedArray((max(n, 1), stride, K), np.uint8)
    ts = sg.PinnedArray((max(n, 1), stride), np.uint32)
    pay = sg.PinnedArray((max(n, 1), stride, P), np.uint8) if want_payload else None
    for a in (counts, keys, ts, pay):
        if a is not None:
            a.array[...] = 0
    sg._check(sg.lib().stage_gpu_scan_batch(snap.h, hk.ptr, sz.ctypes.data, n, stride, read_ts, vis, counts.ptr, keys.ptr,
                                            ts.ptr, pay.ptr if pay else None, sg.IO_HOST))
    out = (counts.array[:n].copy(), keys.array[:n].copy(), ts.array[:n].copy(), pay.array[:n].copy() if pay else None)
    for a in (hk, counts, keys, ts, pay):
        if a is not None:
            a.free()
    return out


def _pinned_probe(snap, keys, read_ts):
    n, K, P = keys.shape[0], snap.schema.key_size, snap.schema.payload_size
    hk = sg.PinnedArray((max(n, 1), K), np.uint8)
    hk.array[:n] = keys
    found = sg.PinnedArray((max(n, 1),), np.uint8)
    ts = sg.PinnedArray((max(n, 1),), np.uint32)
    ko = sg.PinnedArray((max(n, 1), K), np.uint8)
    pay = sg.PinnedArray((max(n, 1), P), np.uint8)
    for a in (found, ts, ko, pay):
        a.array[...] = 0
    sg._check(sg.lib().stage_gpu_probe_batch(snap.h, hk.ptr, n, read_ts, found.ptr, ts.ptr, ko.ptr, pay.ptr, sg.IO_HOST))
    out = (found.array[:n].copy(), ts.array[:n].copy(), ko.array[:n].copy(), pay.array[:n].copy())
    for a in (hk, found, ts, ko, pay):
        a.free()
    return out


@pytest.mark.parametrize("vis", [0, 1])
@pytest.mark.parametrize("scan_size", [10, 100, 1000])
def test_scan_pinned_outputs_match_checker(mvcc_table, scan_size, vis):
    t, snap = mvcc_table
    sk = u32_keys(pyoracle.uniform_keys(20000, 96, seed=5) - 1)
    for read_ts in (0, 1500, 5000):
        want = t.scan(sk, scan_size, read_ts, vis)
        assert_scan_equal(_pinned_scan(snap, sk, scan_size, scan_size, read_ts, vis), want, f"pinned S={scan_size} ts={read_ts}")
        if not ZEROCOPY:
            assert_scan_equal(snap.scan(sk, scan_size, read_ts, vis), want, f"pageable S={scan_size} ts={read_ts}")


def _expand_compact(got, stride):
    """compact result -> the strided arrays the checker returns"""
    counts, keys, ts, pay = got
    n = counts.shape[0]
    K, P = keys.shape[1], pay.shape[1]
    ko, to, po = np.zeros((n, stride, K), np.uint8), np.zeros((n, stride), np.uint32), np.zeros((n, stride, P), np.uint8)
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    for i in range(n):
        c = int(counts[i])
        ko[i, :c], to[i, :c], po[i, :c] = keys[off[i]:off[i] + c], ts[off[i]:off[i] + c], pay[off[i]:off[i] + c]
    return counts, ko, to, po


@pytest.mark.parametrize("scan_size", [10, 1000])
def test_scan_compact_layout_matches_checker(mvcc_table, scan_size):
    """STAGE_IO_COMPACT: rows packed back to back (exactly the rows that exist cross PCIe), several chunks."""
    t, snap = mvcc_table
    n = 300 if scan_size == 1000 else 5000
    sk = u32_keys(pyoracle.uniform_keys(20000, n, seed=11) - 1)
    sizes = np.where(np.arange(n) % 5 == 0, max(scan_size // 3, 1), scan_size).astype(np.uint32)
    for read_ts, vis in ((0, 1), (1500, 1), (0, 0)):
        want = t.scan(sk, sizes, read_ts, vis)
        got = snap.scan_compact(sk, sizes, read_ts, vis)
        assert got[1].shape[0] == int(want[0].sum())
        assert_scan_equal(_expand_compact(got, scan_size), want, f"compact S={scan_size} ts={read_ts} vis={vis}")


def test_compact_with_empty_scans_and_table_end():
    g = sg.Snapshot.generate_ycsb(3000, 1000)
    from stage_indexorganized_b200.ycsb import rank_to_key
    sk = u32_keys(np.concatenate([rank_to_key(3000, np.array([2990, 2999, 0, 1500])), np.array([0x7F7F7F7F], np.uint32)]))
    counts, keys, ts, pay = g.scan_compact(sk, np.array([50, 50, 0, 7, 9], np.uint32), 0, 0)
    assert counts.tolist() == [10, 1, 0, 7, 0] and pay.shape == (18, 1000)
    c2, k2, _, p2 = g.scan(sk, np.array([50, 50, 0, 7, 9], np.uint32), 0, 0)
    assert np.array_equal(c2, counts) and np.array_equal(pay[:10], p2[0, :10]) and np.array_equal(pay[11:18], p2[3, :7])
    assert np.array_equal(keys[10], k2[1, 0])
    g.close()


def test_zero_copy_variant_in_child_process():
    """The same file with STAGE_HOST_PATH=zerocopy: kernels store into the pinned outputs directly."""
    if os.environ.get("STAGE_VARIANT_RUN"):
        pytest.skip("already inside a variant run")
    env = dict(os.environ, STAGE_HOST_PATH="zerocopy", STAGE_VARIANT_RUN="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k", "pinned or registered"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


def test_scan_staged_chunking_many_scans(mvcc_table):
    """More result bytes than one staging chunk (128 MiB): several chunks, both buffers reused."""
    t, snap = mvcc_table
    sk = u32_keys(pyoracle.uniform_keys(20000, 400, seed=9) - 1)
    sizes = np.where(np.arange(400) % 7 == 0, 3, 1000).astype(np.uint32)          # ragged: short scans between long ones
    want = t.scan(sk, sizes, 2000, 1)
    assert_scan_equal(snap.scan(sk, sizes, 2000, 1), want, "pageable, 400 scans")
    assert_scan_equal(_pinned_scan(snap, sk, sizes, 1000, 2000, 1), want, "pinned outputs, 400 scans")


def test_probe_pinned_and_pageable_match_checker(mvcc_table):
    t, snap = mvcc_table
    keys = u32_keys(np.concatenate([pyoracle.uniform_keys(20000, 150000, seed=4), np.array([17, 4242, 70000], np.uint32)]))
    for read_ts in (0, 9, 1500, 0xFFFFFFFF):
        want = t.probe(keys, read_ts)
        assert_probe_equal(_pinned_probe(snap, keys, read_ts), want, f"pinned ts={read_ts}")
        if not ZEROCOPY:
            assert_probe_equal(snap.probe(keys, read_ts), want, f"pageable ts={read_ts}")     # > 1 staging chunk (64 Ki probes)


def test_registered_caller_memory_is_written_in_place(mvcc_table):
    """stage_gpu_host_register: a caller-owned array becomes a zero-copy target."""
    t, snap = mvcc_table
    sk = u32_keys(np.array([5, 9000, 15000], np.uint32))
    n, S = 3, 50
    arena = np.zeros(1 << 20, np.uint8)          # one caller-owned allocation, registered once, carved into the outputs
    counts = arena[0:4 * n].view(np.uint32)
    ts = arena[4096:4096 + 4 * n * S].view(np.uint32).reshape(n, S)
    keys = arena[8192:8192 + 4 * n * S].reshape(n, S, 4)
    pay = arena[16384:16384 + n * S * 1000].reshape(n, S, 1000)
    sizes = np.full(n, S, np.uint32)
    sg._check(sg.lib().stage_gpu_host_register(arena.ctypes.data, arena.nbytes))
    try:
        sg._check(sg.lib().stage_gpu_scan_batch(snap.h, sk.ctypes.data, sizes.ctypes.data, n, S, 3000, 1, counts.ctypes.data,
                                                keys.ctypes.data, ts.ctypes.data, pay.ctypes.data, sg.IO_HOST))
    finally:
        sg._check(sg.lib().stage_gpu_host_unregister(arena.ctypes.data))
    assert_scan_equal((counts, keys, ts, pay), t.scan(sk, S, 3000, 1), "registered")


def test_concurrent_callers_share_one_snapshot(mvcc_table):
    """8 threads, each its own batches of probes and scans on the same handle (contexts instead of a global lock)."""
    t, snap = mvcc_table
    errors = []

    def worker(w):
        try:
            rng = np.random.default_rng(100 + w)
            for it in range(4):
                keys = u32_keys(rng.integers(0, 20100, 5000).astype(np.uint32))
                assert_probe_equal(snap.probe(keys, 2000 + w), t.probe(keys, 2000 + w), f"thread {w} probe {it}")
                sk = u32_keys(rng.integers(0, 20000, 24).astype(np.uint32))
                S = int(rng.choice([7, 200, 1000]))
                assert_scan_equal(_pinned_scan(snap, sk, S, S, 1000 * w, 1), t.scan(sk, S, 1000 * w, 1), f"thread {w} scan {it}")
        except Exception as e:      # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=worker, args=(w,)) for w in range(8)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errors, errors[:3]


def test_slot_order_counter_counts_quirk_scans():
    """The counter sees the scans whose last leaf is truncated in slot order; a consolidated table has none."""
    cls = _checker()
    t, keys, _ = random_table(cls, 3000, 8, 96, seed=3)
    snap = sg.Snapshot.from_tree(t.root, 8, 96)
    snap.counters(reset=True)
    rng = np.random.default_rng(1)
    sk = keys[rng.integers(0, keys.shape[0], 512)]
    assert_scan_equal(snap.scan(sk, 10, 0, 0), t.scan(sk, 10, 0, 0), "tiny rows")
    q = snap.counters().slot_order_scans
    assert q > 0                                   # random insertion order: most leaves are not slot-ordered
    snap.close()
    g = sg.Snapshot.generate_ycsb(5000, 1000)
    g.scan(u32_keys(np.arange(100, dtype=np.uint32)), 10, 0, 0)
    assert g.counters().slot_order_scans == 0
    g.close()


def test_scan_size_above_stride_is_clamped_on_device_and_rejected_on_host(mvcc_table):
    import torch
    t, snap = mvcc_table
    sk = u32_keys(np.array([100, 200], np.uint32))
    with pytest.raises(sg.StageError):
        snap.scan(sk, np.array([30, 10], np.uint32), 0, 0, out_stride=20)
    dev = torch.device("cuda", 0)
    d_sk = torch.from_numpy(sk).to(dev)
    d_sz = torch.tensor([30, 10], dtype=torch.int32, device=dev)
    counts = torch.zeros(2, dtype=torch.int32, device=dev)
    kout = torch.full((3, 20, 4), 0xEE, dtype=torch.uint8, device=dev)       # one guard scan behind the two real ones
    tso = torch.zeros((3, 20), dtype=torch.int32, device=dev)
    pay = torch.full((3, 20, 1000), 0xEE, dtype=torch.uint8, device=dev)
    snap.scan_device(d_sk, d_sz, 2, 20, 0, 0, counts, kout, tso, pay)
    torch.cuda.synchronize()
    assert counts.tolist() == [20, 10]
    want = t.scan(sk, np.array([20, 10], np.uint32), 0, 0)
    assert np.array_equal(pay[:2].cpu().numpy()[0], want[3][0])
    assert int((pay[2] != 0xEE).sum()) == 0 and int((kout[1, 10:] != 0xEE).sum()) == 0    # nothing written past a slab


def test_batching_front_end_coalesces_concurrent_single_operations(mvcc_table):
    """stage_gpu_batcher_*: 16 threads, each one operation at a time with its OWN read id (the reference's driver shape:
    one executor per operation on N worker threads, ycsb_mixed.cpp:83-93). Every answer must be the checker's answer
    for that (key, read id); the operations must have been coalesced (fewer batches than operations)."""
    t, snap = mvcc_table
    b = sg.Batcher(snap, max_batch=64, max_wait_us=200)
    errors = []
    n_threads, per = 16, 150

    def worker(w):
        try:
            rng = np.random.default_rng(500 + w)
            keys = rng.integers(0, 20050, per).astype(np.uint32)
            stamps = rng.integers(0, 3200, per)
            for k, ts in zip(keys, stamps):
                kk = u32_keys(np.array([k], np.uint32))
                if k % 5 == 0:
                    S = int(rng.choice([5, 40, 300]))
                    vis = int(k % 2)
                    c, ko, to, po = b.scan(kk[0], S, int(ts), vis)
                    wc, wk, wt, wp = t.scan(kk, S, int(ts), vis)
                    assert c == wc[0] and np.array_equal(ko[:c], wk[0, :c]) and np.array_equal(to[:c], wt[0, :c]) \
                        and np.array_equal(po[:c], wp[0, :c]), f"thread {w} scan key {k} ts {ts}"
                else:
                    f, s, ko, po = b.probe(kk[0], int(ts))
                    wf, ws, wk, wp = t.probe(kk, int(ts))
                    assert f == wf[0], f"thread {w} probe key {k} ts {ts}"
                    if f:
                        assert s == ws[0] and np.array_equal(ko, wk[0]) and np.array_equal(po, wp[0]), f"thread {w} key {k}"
        except Exception as e:      # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=worker, args=(w,)) for w in range(n_threads)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    batches, ops = b.stats()
    b.close()
    assert not errors, errors[:3]
    assert ops == n_threads * per and batches < ops, (batches, ops)


@pytest.mark.skipif(ZEROCOPY, reason="the small-call path belongs to the default host path")
def test_small_call_path_and_its_size_boundary(mvcc_table):
    """Calls whose keys + results fit the 192 KiB pinned slab of the call context run without staging copies
    (probe_host_small / scan_host_small, csrc/stage_gpu.cu); one operation more and the staged path takes over. Both sides
    of the boundary, both layouts, pageable buffers pre-filled with a pattern: same answers as the checker, rows of
    missed probes and slots beyond row_counts untouched on the small path."""
    t, snap = mvcc_table
    K, P = 4, 1000
    per_probe = K + 1 + 4 + K + P + 64                     # generous: carve-outs are 16-byte aligned
    n_small = (192 << 10) // per_probe - 2
    rng = np.random.default_rng(91)
    for n in (1, 3, n_small, n_small + 8, 4 * n_small):
        keys = u32_keys(rng.integers(0, 20050, n).astype(np.uint32))
        keys[0] = u32_keys(np.array([20040], np.uint32))[0]                      # a miss in every batch
        for read_ts in (0, 1500, 5000):
            want = t.probe(keys, read_ts)
            found = np.full(n, 7, np.uint8)
            ts = np.zeros(n, np.uint32)
            ko = np.full((n, K), 0xEE, np.uint8)
            po = np.full((n, P), 0xEE, np.uint8)
            sg._check(sg.lib().stage_gpu_probe_batch(snap.h, keys.ctypes.data, n, read_ts, found.ctypes.data, ts.ctypes.data,
                                                     ko.ctypes.data, po.ctypes.data, sg.IO_HOST))
            hit = want[0] == 1
            assert np.array_equal(found, want[0]) and np.array_equal(ts[hit], want[1][hit])
            assert np.array_equal(ko[hit], want[2][hit]) and np.array_equal(po[hit], want[3][hit]), (n, read_ts)
            if n <= n_small:
                assert (po[~hit] == 0xEE).all() and (ko[~hit] == 0xEE).all(), "small path leaves missed rows untouched"
    # scans: 1 x 100 rows (102 KB of results: small), 2 x 100 rows (staged), 12 x 10 rows (small), strided and compact
    for n, S in ((1, 100), (2, 100), (12, 10), (1, 1), (40, 10)):
        sk = u32_keys(rng.integers(0, 20000, n).astype(np.uint32))
        sizes = np.full(n, S, np.uint32)
        for vis in (0, 1):
            wc, wk, wt, wp = t.scan(sk, S, 1500, vis)
            for compact in (False, True):
                counts = np.zeros(n, np.uint32)
                ko = np.full((n * S, K), 0xEE, np.uint8)
                to = np.full(n * S, 0xEEEEEEEE, np.uint32)
                po = np.full((n * S, P), 0xEE, np.uint8)
                sg._check(sg.lib().stage_gpu_scan_batch(snap.h, sk.ctypes.data, sizes.ctypes.data, n, S, 1500, vis, counts.ctypes.data,
                                                        ko.ctypes.data, to.ctypes.data, po.ctypes.data,
                                                        sg.IO_HOST | (sg.IO_COMPACT if compact else 0)))
                assert np.array_equal(counts, wc), (n, S, vis, compact)
                base = 0
                for i in range(n):
                    c = int(wc[i])
                    o = base if compact else i * S
                    assert np.array_equal(ko[o:o + c], wk[i, :c]) and np.array_equal(to[o:o + c], wt[i, :c]) \
                        and np.array_equal(po[o:o + c], wp[i, :c]), (n, S, vis, compact, i)
                    base += c
