"""GPU: structural refresh (stage_gpu_snapshot_refresh_from_tree) — after committed inserts (with leaf splits), updates
and deletes on the host tree, the refreshed snapshot must answer exactly like the tree (and like a snapshot flattened
from scratch), while only the leaves that changed are staged again (the rest is carried over on the device)."""
import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from helpers import assert_probe_equal, assert_scan_equal, u32_keys, ycsb_int_table, random_table
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def checker_classes():
    return [pyoracle.OracleTable] + ([pyoracle.RefTable] if pyoracle.have_ref() else [])


def _check(snap, t, probes, starts, S, stamps, what):
    for read_ts in stamps:
        assert_probe_equal(snap.probe(probes, read_ts), t.probe(probes, read_ts), f"{what} probe ts={read_ts}")
        for vis in (0, 1):
            assert_scan_equal(snap.scan(starts, S, read_ts, vis), t.scan(starts, S, read_ts, vis),
                              f"{what} scan ts={read_ts} vis={vis}")


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
def test_refresh_after_updates_inserts_deletes(cls):
    N, P = 40000, 200
    t = ycsb_int_table(cls, N, P)
    rng = np.random.default_rng(31)
    assert t.commit_update_batch(u32_keys(rng.integers(0, N, 3000)), 10) == 3000      # chains before the first snapshot
    snap = sg.Snapshot.from_tree(t.root, 4, P)
    probes = u32_keys(np.concatenate([rng.integers(0, N + 50, 6000), 0x10000007 + 256 * rng.integers(0, 2600, 3000)]))
    starts = np.concatenate([u32_keys(rng.integers(0, N, 96)), u32_keys(0x10000007 + 256 * rng.integers(0, 2500, 32))])
    stamps = (0, 10, 1500, 3009, 5000, 5010, 5020, 6000, 6500, 9000)
    try:
        _check(snap, t, probes, starts, 150, (0, 1500, 9000), "initial")
        L0 = snap.info().n_leaves

        # round 1: a few committed updates -> only their leaves are re-transformed
        k = rng.integers(0, N, 12).astype(np.uint32)
        assert t.commit_update_batch(u32_keys(k), 5000) == 12
        s1, st = snap.refresh_from_tree(t.root)
        assert st.leaves_total == L0 and 1 <= st.leaves_transformed <= 12
        assert st.leaves_carried == L0 - st.leaves_transformed and st.versions_added >= 12
        assert st.host_bytes_staged < 0.25 * L0 * 65536
        _check(s1, t, probes, starts, 150, stamps, "after updates")
        _check(snap, t, probes[:0], starts[:0], 1, (), "")
        snap.close()
        snap = s1

        # round 2: 2500 inserts into one region of the key order (same low byte) -> leaf splits there
        ik = (0x10000007 + 256 * np.arange(2500)).astype(np.uint32)
        rng.shuffle(ik)
        ipay = np.repeat((ik >> 8).astype(np.uint8)[:, None], P, axis=1)
        assert t.insert(u32_keys(ik), ipay, 5010) == 2500
        s2, st = snap.refresh_from_tree(t.root)
        assert st.leaves_total > L0, "the inserts split leaves"
        assert st.rows == N + 2500
        assert st.leaves_carried >= L0 - 8 and st.leaves_transformed <= 8 + 2 * (st.leaves_total - L0)
        _check(s2, t, probes, starts, 150, stamps, "after inserts")
        snap.close()
        snap = s2

        # round 3: deletes (old and freshly inserted rows) + updates of inserted rows
        dk = np.concatenate([rng.choice(N, 40, replace=False).astype(np.uint32), ik[:200]])
        for key in u32_keys(dk):
            t.delete(key)
        assert t.commit_update_batch(u32_keys(ik[300:400]), 6000) == 100
        s3, st = snap.refresh_from_tree(t.root)
        assert st.rows == N + 2500 - 240
        assert 0 < st.leaves_transformed < st.leaves_total // 2
        _check(s3, t, probes, starts, 150, stamps, "after deletes")
        snap.close()
        snap = s3

        # the refreshed snapshot equals one flattened from scratch (same rows, leaves and answers) ...
        with sg.Snapshot.from_tree(t.root, 4, P) as fresh:
            a, b = snap.info(), fresh.info()
            assert (a.n_rows, a.n_leaves, a.max_leaf_rows) == (b.n_rows, b.n_leaves, b.max_leaf_rows)
            for read_ts in (0, 5005, 9000):
                assert_probe_equal(snap.probe(probes, read_ts), fresh.probe(probes, read_ts), "vs fresh")
        # ... and a refresh with nothing changed carries every leaf over
        s4, st = snap.refresh_from_tree(t.root)
        assert st.leaves_transformed == 0 and st.host_bytes_staged == 0 and st.versions_added == 0
        _check(s4, t, probes, starts, 150, (0, 5015, 9000), "no-op refresh")
        s4.close()
    finally:
        snap.close()


def test_refresh_string_keys_random_order_inserts():
    """16-byte keys (memcmp order, 4 key words), inserts in random order everywhere: most leaves change and the
    refresh degenerates to a rebuild — still exact."""
    t, keys, pay = random_table(pyoracle.OracleTable, 6000, 16, 120, seed=8)
    rng = np.random.default_rng(2)
    with sg.Snapshot.from_tree(t.root, 16, 120) as snap:
        more = rng.integers(0, 256, size=(3000, 16), dtype=np.uint8)
        mpay = np.repeat(more[:, :1], 120, axis=1)
        n_new = t.insert(more, mpay, 7)
        assert n_new >= 2990
        for k in keys[:150]:
            t.delete(k)
        s2, st = snap.refresh_from_tree(t.root)
        try:
            assert st.rows == keys.shape[0] + n_new - 150
            probes = np.concatenate([keys[:2000], more[:1000], rng.integers(0, 256, size=(500, 16), dtype=np.uint8)])
            for read_ts in (0, 6, 7, 100):
                assert_probe_equal(s2.probe(probes, read_ts), t.probe(probes, read_ts), f"ts={read_ts}")
                assert_scan_equal(s2.scan(probes[:80], 90, read_ts, 1), t.scan(probes[:80], 90, read_ts, 1), "scan")
        finally:
            s2.close()


def test_refresh_needs_a_tree_snapshot():
    with sg.Snapshot.generate_ycsb(1000, 64) as snap:
        t = ycsb_int_table(pyoracle.OracleTable, 1000, 64)
        with pytest.raises(sg.StageError):
            snap.refresh_from_tree(t.root)              # generated snapshots carry no host fingerprints


# ---- snapshots with writers in flight --------------------------------------------------------
@pytest.mark.skipif(not pyoracle.have_ref(), reason="needs the live reference (EphemeralPool undo copies)")
def test_snapshot_with_updates_in_flight_reads_the_undo_copies():
    """Updates begun through the reference's LeafNode::Update and NOT finalized: BTree::Read serves the undo copy
    (old payload, old commit id, saved chain). stage_gpu_snapshot_from_tree_live stages exactly that; the plain
    from_tree refuses such a tree."""
    N, P = 8000, 200
    t = ycsb_int_table(pyoracle.RefTable, N, P)
    rng = np.random.default_rng(44)
    assert t.commit_update_batch(u32_keys(rng.integers(0, N, 1500)), 10) == 1500        # committed history first
    assert t.commit_update_batch(u32_keys(rng.integers(0, N, 1500)), 2000) == 1500
    probes = u32_keys(rng.integers(0, N + 20, 8000))
    stamps = (0, 10, 800, 1509, 2000, 2900, 3499, 5000, 9000)
    before = {ts: t.probe(probes, ts) for ts in stamps}
    sk = u32_keys(rng.integers(0, N, 48))
    scans_before = {ts: t.scan(sk, 120, ts, 1) for ts in stamps}
    # 300 updates in flight, half of them on rows that already have a committed chain
    flying = rng.choice(N, 300, replace=False).astype(np.uint32)
    flying_before = t.probe(u32_keys(flying), 9000)
    quiesced = sg.Snapshot.from_tree(t.root, 4, P)         # taken before the writers start
    for i, k in enumerate(flying):                         # new bytes differ from every committed byte of the row
        assert t.begin_update(u32_keys([k])[0], np.full(P, (int(k) & 0xFF) ^ 0x5A, np.uint8), 7000 + i) == 0
    with pytest.raises(sg.StageError):
        sg.Snapshot.from_tree(t.root, 4, P)                                            # not quiesced
    fn, ctx = t.undo_resolver()
    with sg.Snapshot.from_tree_live(t.root, 4, P, fn, ctx) as snap:
        assert snap.info().n_rows == N
        for ts in stamps:
            now = t.probe(probes, ts)                      # the reference, writers still in flight
            assert_probe_equal(snap.probe(probes, ts), now, f"in-flight probe ts={ts}")
            assert_probe_equal(now, before[ts], f"reference serves the committed state ts={ts}")
            # scans: the snapshot holds the committed rows (the reference's iterator would read dirty bytes)
            assert_scan_equal(snap.scan(sk, 120, ts, 1), scans_before[ts], f"in-flight scan ts={ts}")
        assert_probe_equal(snap.probe(u32_keys(flying), 9000), flying_before, "no uncommitted bytes on the device")
    # refreshing the earlier, quiesced snapshot while the writers are in flight: the leaves they touched are staged
    # again (their metadata changed) from the undo copies, everything else is carried over; same answers
    with quiesced:
        with pytest.raises(sg.StageError):
            quiesced.refresh_from_tree(t.root)
        nxt, st = quiesced.refresh_from_tree(t.root, fn, ctx)
        with nxt:
            assert st.rows == N and 0 < st.leaves_transformed <= st.leaves_total
            for ts in stamps:
                assert_probe_equal(nxt.probe(probes, ts), before[ts], f"live refresh probe ts={ts}")
                assert_scan_equal(nxt.scan(sk, 120, ts, 1), scans_before[ts], f"live refresh scan ts={ts}")


def _snapshot_is_consistent(snap, N, P, keys, first_cid, committed):
    """What a snapshot taken while the writer ran must satisfy, whatever instant it caught (the writer's transactions are
    known: update i rewrites column [100,200) of keys[i] with (uint8)(first_cid + i) and commits with id first_cid + i):
      * every row is there; its in-leaf commit id is 0 or the id of one of ITS OWN committed updates;
      * the row image belongs to that version: column = (uint8) commit id (or the loader's byte), the rest untouched —
        never a new payload under an old commit id, never a torn row;
      * the version chain is intact: a reader just older than the in-leaf version gets the PREVIOUS update of that key
        (or, before the second update, nothing: the loader's version ends the walk, executor.h:418)."""
    allk = np.arange(N, dtype=np.uint32)
    f, ts, _, pay = snap.probe(u32_keys(allk), 0xFFFFFFFF)
    assert f.all(), "rows missing"
    base = (allk & 0xFF).astype(np.uint8)
    col = np.where(ts > 0, (ts & 0xFF).astype(np.uint8), base)
    assert (pay[:, :100] == base[:, None]).all() and (pay[:, 200:] == base[:, None]).all(), "bytes outside the column changed"
    assert (pay[:, 100:200] == col[:, None]).all(), "column does not belong to the commit id of the row"
    cids = first_cid + np.nonzero(committed)[0]
    ck = keys[np.nonzero(committed)[0]]
    order = np.lexsort((cids, ck))
    ck, cids = ck[order], cids[order]
    upd = ts > 0
    # the in-leaf commit id is one of the key's own updates
    pos = np.searchsorted(ck.astype(np.uint64) << np.uint64(32) | cids.astype(np.uint64),
                          allk[upd].astype(np.uint64) << np.uint64(32) | ts[upd].astype(np.uint64))
    assert (pos < len(ck)).all() and (ck[pos] == allk[upd]).all() and (cids[pos] == ts[upd]).all(), "foreign commit id"
    # previous version through the chain
    has_prev = (pos > 0) & (ck[np.maximum(pos - 1, 0)] == allk[upd])
    prev_cid = np.where(has_prev, cids[np.maximum(pos - 1, 0)], 0)
    uk, uts = allk[upd], ts[upd]
    for lo in range(0, len(uk), 4096):                       # one read id per probe: small batches per distinct stamp
        sl = slice(lo, lo + 4096)
        for k_, t_, p_, h_ in zip(uk[sl][:64], uts[sl][:64], prev_cid[sl][:64], has_prev[sl][:64]):
            f1, s1, _, p1 = snap.probe(u32_keys(np.array([k_], np.uint32)), int(t_) - 1)
            if h_:
                assert f1[0] == 1 and s1[0] == p_ and p1[0, 150] == (p_ & 0xFF), "version chain lost a version"
            else:
                assert f1[0] == 0, "a version older than the first update must not be visible"


@pytest.mark.skipif(not pyoracle.have_ref(), reason="needs the unmodified reference (oracle/_ref)")
def test_snapshots_while_update_transactions_are_running():
    """from_tree_live / refresh_from_tree_live with a writer thread that is actually running update
    transactions (LeafNode::Update + commit install, the reference's own steps and order). Every leaf copy is validated
    against the discovery pass and re-done when a writer got in between; the snapshots must be consistent."""
    N, P = 30000, 1000
    t = ycsb_int_table(pyoracle.RefTable, N, P)
    rng = np.random.default_rng(5)
    keys = rng.integers(0, N, 3_000_000).astype(np.uint32)
    fn, ctx = t.undo_resolver()
    w = pyoracle.Writer(t, keys, 100)
    snaps = []
    try:
        first = None
        for it in range(4):
            assert w.alive(), "the writer finished before the snapshots were taken: enlarge the key stream"
            if first is None:
                first = sg.Snapshot.from_tree_live(t.root, 4, P, fn, ctx)
                snaps.append(first)
            else:
                nxt, st = snaps[-1].refresh_from_tree(t.root, fn, ctx)
                assert st.leaves_transformed > 0
                snaps.append(nxt)
            snaps.append(sg.Snapshot.from_tree_live(t.root, 4, P, fn, ctx))
    finally:
        tried, committed = w.stop()
    assert tried > 10000 and committed[:tried].sum() > 0.9 * tried
    for s in snaps:
        _snapshot_is_consistent(s, N, P, keys[:tried], 100, committed[:tried])
        s.close()
    # quiesced again: the plain path agrees with the tree
    with sg.Snapshot.from_tree(t.root, 4, P) as q:
        pk = u32_keys(rng.integers(0, N, 20000))
        for read_ts in (0, 50000, 0xFFFFFFFF):
            assert_probe_equal(q.probe(pk, read_ts), t.probe(pk, read_ts), f"quiesced ts={read_ts}")


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
def test_refresh_in_place_shifts_rows_inside_the_payload_buffer(cls):
    """stage_gpu_snapshot_refresh_in_place: the same rounds (updates with new chains, inserts with leaf splits in the
    middle of the key order -> rows to the right shift right; deletes -> rows shift left; nothing changed) with the
    payload rows moved inside the existing buffer. Answers equal the tree's and a from-scratch snapshot's, the device
    footprint does not grow beyond the small columns, and a snapshot without room refuses and stays usable."""
    N, P = 40000, 200
    t = ycsb_int_table(cls, N, P)
    rng = np.random.default_rng(77)
    assert t.commit_update_batch(u32_keys(rng.integers(0, N, 3000)), 10) == 3000
    probes = u32_keys(np.concatenate([rng.integers(0, N + 50, 6000), 0x10000007 + 256 * rng.integers(0, 2600, 3000)]))
    starts = np.concatenate([u32_keys(rng.integers(0, N, 96)), u32_keys(0x10000007 + 256 * rng.integers(0, 2500, 32))])
    stamps = (0, 10, 1500, 3009, 5000, 5010, 5020, 6000, 6500, 9000)
    ik = (0x10000007 + 256 * np.arange(2500)).astype(np.uint32)
    rng.shuffle(ik)
    ipay = np.repeat((ik >> 8).astype(np.uint8)[:, None], P, axis=1)

    tight = sg.Snapshot.from_tree(t.root, 4, P)                       # no slack: inserts cannot be taken in place
    sg.set_row_slack(0.10)
    try:
        snap = sg.Snapshot.from_tree(t.root, 4, P)
    finally:
        sg.set_row_slack(0.0)
    try:
        bytes0 = snap.info().device_bytes
        L0 = snap.info().n_leaves
        # round 1: committed updates (new chains are appended to the version table)
        assert t.commit_update_batch(u32_keys(rng.integers(0, N, 12).astype(np.uint32)), 5000) == 12
        st = snap.refresh_in_place(t.root)
        assert st.leaves_total == L0 and 1 <= st.leaves_transformed <= 12 and st.versions_added >= 12
        _check(snap, t, probes, starts, 150, stamps, "in place, after updates")
        # round 2: inserts into one region -> splits; every row behind them shifts right
        assert t.insert(u32_keys(ik), ipay, 5010) == 2500
        with pytest.raises(sg.StageError):
            tight.refresh_in_place(t.root)                            # no room: refused before anything moved
        assert tight.info().n_rows == N and tight.probe(probes[:16], 0)[0].shape == (16,)     # untouched, still answers
        st = snap.refresh_in_place(t.root)
        assert st.leaves_total > L0 and st.rows == N + 2500
        assert st.leaves_carried >= L0 - 8
        _check(snap, t, probes, starts, 150, stamps, "in place, after inserts")
        # round 3: deletes everywhere (rows shift left) + updates of inserted rows
        dk = np.concatenate([rng.choice(N, 400, replace=False).astype(np.uint32), ik[:200]])
        for key in u32_keys(dk):
            t.delete(key)
        assert t.commit_update_batch(u32_keys(ik[300:400]), 6000) == 100
        st = snap.refresh_in_place(t.root)
        assert st.rows == N + 2500 - 600
        _check(snap, t, probes, starts, 150, stamps, "in place, after deletes")
        # round 4: nothing changed -> nothing staged, the version table is kept as it is
        st = snap.refresh_in_place(t.root)
        assert st.leaves_transformed == 0 and st.host_bytes_staged == 0 and st.versions_added == 0
        _check(snap, t, probes, starts, 150, (0, 5015, 9000), "in place, no-op")
        with sg.Snapshot.from_tree(t.root, 4, P) as fresh:
            a, b = snap.info(), fresh.info()
            assert (a.n_rows, a.n_leaves, a.max_leaf_rows) == (b.n_rows, b.n_leaves, b.max_leaf_rows)
            for read_ts in (0, 5005, 9000):
                assert_probe_equal(snap.probe(probes, read_ts), fresh.probe(probes, read_ts), "vs fresh")
                assert_scan_equal(snap.scan(starts, 150, read_ts, 1), fresh.scan(starts, 150, read_ts, 1), "vs fresh")
        # the payload buffer was never duplicated: the footprint stays within the small columns + version growth
        assert snap.info().device_bytes < bytes0 * 1.15
        # and the ordinary refresh still works on a snapshot that went through in-place rounds
        assert t.commit_update_batch(u32_keys(ik[500:520]), 7000) == 20
        s2, st = snap.refresh_from_tree(t.root)
        _check(s2, t, probes, starts, 150, (0, 6999, 7010, 9000), "two-table refresh after in-place rounds")
        s2.close()
    finally:
        snap.close()
        tight.close()


def test_refresh_in_place_large_shifts():
    """A 250 MB payload column whose rows shift by less than the 64 MiB bounce buffer (overlapping runs, several chunks,
    right then left) and by more (chunks that never reach into their own source): keys with low byte 0x80 sort first in
    the reference order, so inserting / deleting them moves nearly every row of the table."""
    N, P = 250_000, 1000
    t = ycsb_int_table(pyoracle.OracleTable, N, P)
    rng = np.random.default_rng(5)
    sg.set_row_slack(0.40)
    try:
        snap = sg.Snapshot.from_tree(t.root, 4, P)
    finally:
        sg.set_row_slack(0.0)

    def front_keys(lo, n):
        return (0x80 + 256 * (1000 + lo + np.arange(n))).astype(np.uint32)

    def check(what, extra):
        probes = u32_keys(np.concatenate([rng.integers(0, N, 20000).astype(np.uint32), extra]))
        starts = u32_keys(np.concatenate([rng.integers(0, N, 48).astype(np.uint32), extra[:16]]))
        for read_ts in (0, 50):
            assert_probe_equal(snap.probe(probes, read_ts), t.probe(probes, read_ts), what)
            assert_scan_equal(snap.scan(starts, 300, read_ts, 1), t.scan(starts, 300, read_ts, 1), what)

    try:
        bytes0 = snap.info().device_bytes
        a = front_keys(0, 3000)
        assert t.insert(u32_keys(a), np.repeat((a >> 8).astype(np.uint8)[:, None], P, axis=1), 20) == 3000
        st = snap.refresh_in_place(t.root)                    # ~3 MB to the right, through the bounce buffer
        assert st.rows == N + 3000 and st.leaves_carried > 0.9 * st.leaves_total
        check("right shift, bounce", a)
        for key in u32_keys(a[:2000]):
            t.delete(key)
        st = snap.refresh_in_place(t.root)                    # ~2 MB to the left
        assert st.rows == N + 1000
        check("left shift, bounce", a)
        b = front_keys(3000, 70_000)
        assert t.insert(u32_keys(b), np.repeat((b >> 8).astype(np.uint8)[:, None], P, axis=1), 30) == 70_000
        st = snap.refresh_in_place(t.root)                    # 70 MB to the right: direct chunks
        assert st.rows == N + 71_000
        check("right shift, direct chunks", np.concatenate([a, b[:5000]]))
        for key in u32_keys(b[:68_000]):
            t.delete(key)
        st = snap.refresh_in_place(t.root)                    # 68 MB to the left
        assert st.rows == N + 3000
        check("left shift, direct chunks", np.concatenate([a, b[-5000:]]))
        with sg.Snapshot.from_tree(t.root, 4, P) as fresh:
            probes = u32_keys(rng.integers(0, N, 50000).astype(np.uint32))
            assert_probe_equal(snap.probe(probes, 99), fresh.probe(probes, 99), "vs fresh")
        assert snap.info().device_bytes < bytes0 * 1.1
    finally:
        snap.close()
        t.close()


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
def test_version_table_compaction_after_refreshes(cls):
    """Every refresh stages the chains of the leaves it re-transforms again; the previous copies stay behind. After a
    few rounds stage_gpu_snapshot_compact_versions must drop exactly the unreachable ones: answers at every read id are
    the tree's before and after, the table shrinks to what a from-scratch snapshot holds, a second call drops nothing."""
    N, P = 30000, 120
    t = ycsb_int_table(cls, N, P)
    rng = np.random.default_rng(41)
    hot = rng.integers(0, N, 400).astype(np.uint32)                    # the same rows again and again: long chains
    assert t.commit_update_batch(u32_keys(rng.integers(0, N, 4000)), 10) == 4000
    snap = sg.Snapshot.from_tree(t.root, 4, P)
    probes = u32_keys(np.concatenate([rng.integers(0, N + 20, 8000).astype(np.uint32), hot]))
    starts = u32_keys(rng.integers(0, N, 64))
    stamps = [0, 10, 2000, 4009]
    try:
        assert snap.compact_versions() == (snap.info().n_versions, 0)                  # fresh: everything is reachable
        cid = 5000
        for rnd in range(4):
            ks = np.concatenate([hot[rng.integers(0, len(hot), 300)], rng.integers(0, N, 200).astype(np.uint32)])
            assert t.commit_update_batch(u32_keys(ks), cid) == len(ks)
            stamps += [cid, cid + 250, cid + 499]
            cid += 1000
            if rnd % 2:
                snap.refresh_in_place(t.root)
            else:
                s2, _ = snap.refresh_from_tree(t.root)
                snap.close()
                snap = s2
        stamps.append(cid + 5)
        before = snap.info().n_versions
        _check(snap, t, probes, starts, 120, stamps, "before compaction")
        kept, dropped = snap.compact_versions()
        with sg.Snapshot.from_tree(t.root, 4, P) as fresh:
            assert kept == fresh.info().n_versions, "what a from-scratch snapshot stages is what stays"
        assert dropped > 0 and kept + dropped == before and snap.info().n_versions == kept
        _check(snap, t, probes, starts, 120, stamps, "after compaction")
        assert snap.compact_versions() == (kept, 0)
        # the compacted table keeps working as the base of further refreshes and device-side updates
        assert t.commit_update_batch(u32_keys(hot[:50]), cid + 100) == 50
        s2, st = snap.refresh_from_tree(t.root)
        snap.close()
        snap = s2
        _check(snap, t, probes, starts, 120, stamps + [cid + 120, cid + 200], "refresh after compaction")
    finally:
        snap.close()
        t.close()
