"""GPU: incremental refresh (stage_gpu_snapshot_apply_updates, csrc/refresh.cuh) — committed updates applied on the
device must leave exactly the state the reference's commit leaves in the tree
(src/concurrency/transaction_manager.cpp:610-676): checked by flattening the SAME tree before the updates, applying
them on the GPU, and comparing probes / scans at timestamps straddling the commit ids with the checker's tree that
took the updates on the host."""
import numpy as np
import pytest
import torch

import stage_indexorganized_b200 as sg
from helpers import assert_probe_equal, assert_scan_equal, u32_keys, ycsb_int_table
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def checker_classes():
    return [pyoracle.OracleTable] + ([pyoracle.RefTable] if pyoracle.have_ref() else [])


def ycsb_patch(P, cids):
    """The update the checker's commit_update_batch applies: payload bytes [lo, hi) <- (uint8) commit id."""
    lo, hi = (100 if P > 100 else 0), (200 if P > 200 else P)
    return lo, np.repeat((cids & 0xFF).astype(np.uint8)[:, None], hi - lo, axis=1)


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
@pytest.mark.parametrize("P,device", [(1000, False), (64, False), (1000, True)])
def test_apply_updates_matches_host_commits(cls, P, device):
    N = 6000
    t = ycsb_int_table(cls, N, P)
    rng = np.random.default_rng(21)
    snap = sg.Snapshot.from_tree(t.root, 4, P)            # flattened BEFORE the updates
    try:
        cid = 10
        for batch in range(3):                            # several batches: version table growth, chains across calls
            n = (1500, 700, 1)[batch]
            k = rng.integers(0, N + 40, n).astype(np.uint32)          # repeats within a batch, some absent keys
            if batch == 1:
                k[:50] = k[50:100]                        # guaranteed duplicates
            keys = u32_keys(k)
            cids = np.arange(cid, cid + n, dtype=np.uint32)
            off, patches = ycsb_patch(P, cids)
            present = int((k < N).sum())
            assert t.commit_update_batch(keys, cid) == present
            if device:
                got = snap.apply_updates(torch.from_numpy(keys).cuda(), torch.from_numpy(patches).cuda(),
                                         torch.from_numpy(cids.view(np.int32)).cuda(), off, device=True)
            else:
                got = snap.apply_updates(keys, patches, cids, off)
            assert got == present
            cid += n
        assert snap.info().n_versions == 2201
        probes = u32_keys(rng.integers(0, N + 10, 6000))
        sk = u32_keys(rng.integers(0, N, 100))
        for read_ts in (0, 9, 10, 800, 1509, 1510, 2000, 2210, 5000):
            assert_probe_equal(snap.probe(probes, read_ts), t.probe(probes, read_ts), f"ts={read_ts}")
            for vis in (0, 1):
                assert_scan_equal(snap.scan(sk, 150, read_ts, vis), t.scan(sk, 150, read_ts, vis), f"ts={read_ts} vis={vis}")
    finally:
        snap.close()


def test_apply_updates_on_top_of_staged_chains():
    """Snapshot taken when chains already exist, then refreshed: old and new versions live in one version table."""
    N, P = 4000, 200
    t = ycsb_int_table(pyoracle.OracleTable, N, P)
    rng = np.random.default_rng(5)
    first = u32_keys(rng.integers(0, N, 900))
    assert t.commit_update_batch(first, 10) == 900
    with sg.Snapshot.from_tree(t.root, 4, P) as snap:
        v0 = snap.info().n_versions
        k = rng.integers(0, N, 1200).astype(np.uint32)
        cids = np.arange(2000, 3200, dtype=np.uint32)
        off, patches = ycsb_patch(P, cids)
        assert t.commit_update_batch(u32_keys(k), 2000) == 1200
        assert snap.apply_updates(u32_keys(k), patches, cids, off) == 1200
        assert snap.info().n_versions == v0 + 1200
        probes = u32_keys(rng.integers(0, N, 5000))
        for read_ts in (0, 10, 500, 909, 1999, 2000, 2600, 3199, 9999):
            assert_probe_equal(snap.probe(probes, read_ts), t.probe(probes, read_ts), f"ts={read_ts}")
            assert_scan_equal(snap.scan(probes[:64], 100, read_ts, 1), t.scan(probes[:64], 100, read_ts, 1), f"scan ts={read_ts}")


def test_apply_updates_on_generated_table():
    """Device-built table + device-applied updates == host-loaded table + host commits (probes)."""
    N, P = 50000, 1000
    t = ycsb_int_table(pyoracle.OracleTable, N, P)
    rng = np.random.default_rng(9)
    k = rng.integers(0, N, 8000).astype(np.uint32)
    cids = np.arange(10, 8010, dtype=np.uint32)
    off, patches = ycsb_patch(P, cids)
    assert t.commit_update_batch(u32_keys(k), 10) == 8000
    with sg.Snapshot.generate_ycsb(N, P) as snap:
        assert snap.apply_updates(u32_keys(k), patches, cids, off) == 8000
        probes = u32_keys(rng.integers(0, N + 5, 20000))
        for read_ts in (0, 10, 4000, 8009, 20000):
            assert_probe_equal(snap.probe(probes, read_ts), t.probe(probes, read_ts), f"ts={read_ts}")


def test_apply_updates_argument_errors():
    with sg.Snapshot.generate_ycsb(100, 64) as snap:
        k = u32_keys(np.arange(4))
        with pytest.raises(sg.StageError):
            snap.apply_updates(k, np.zeros((4, 65), np.uint8), np.arange(4, dtype=np.uint32))       # patch > payload
        with pytest.raises(sg.StageError):
            snap.apply_updates(k, np.zeros((4, 8), np.uint8), np.arange(4, dtype=np.uint32), 60)    # runs past the row
        assert snap.apply_updates(k[:0], np.zeros((0, 8), np.uint8), np.zeros(0, np.uint32)) == 0


def test_snapshot_save_load_roundtrip(tmp_path):
    """stage_gpu_snapshot_save / load: a restored snapshot answers every read bit-identically (version chains,
    string keys, a refreshed table), and can keep taking updates."""
    N, P = 5000, 200
    t = ycsb_int_table(pyoracle.OracleTable, N, P)
    rng = np.random.default_rng(13)
    assert t.commit_update_batch(u32_keys(rng.integers(0, N, 800)), 10) == 800
    for k in rng.choice(N, 100, replace=False):
        t.delete(u32_keys([k])[0])
    path = str(tmp_path / "snap.bin")
    probes = u32_keys(rng.integers(0, N + 10, 4000))
    with sg.Snapshot.from_tree(t.root, 4, P) as snap:
        snap.save(path)
        want_info = snap.info()
    with sg.Snapshot.load(path) as back:
        i = back.info()
        assert (i.n_rows, i.n_leaves, i.n_versions, i.max_leaf_rows) == (want_info.n_rows, want_info.n_leaves,
                                                                        want_info.n_versions, want_info.max_leaf_rows)
        assert back.schema.key_size == 4 and back.schema.payload_size == P
        for read_ts in (0, 10, 400, 809, 5000):
            assert_probe_equal(back.probe(probes, read_ts), t.probe(probes, read_ts), f"ts={read_ts}")
            for vis in (0, 1):
                assert_scan_equal(back.scan(probes[:64], 120, read_ts, vis), t.scan(probes[:64], 120, read_ts, vis), "scan")
        k = rng.integers(0, N, 300).astype(np.uint32)
        cids = np.arange(3000, 3300, dtype=np.uint32)
        off, patches = ycsb_patch(P, cids)
        present = t.commit_update_batch(u32_keys(k), 3000)
        assert back.apply_updates(u32_keys(k), patches, cids, off) == present
        assert_probe_equal(back.probe(probes, 3100), t.probe(probes, 3100), "after refresh of a loaded snapshot")
    with open(path, "r+b") as f:
        f.write(b"XXXXXXXX")
    with pytest.raises(sg.StageError):
        sg.Snapshot.load(path)
