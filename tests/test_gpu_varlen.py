"""GPU: tables whose keys differ in length (BaseNode::KeyCompare's "prefix, then length" rule, b_tree.h:116-134) —
the shape of test/testing_btree.cpp's fixtures (std::to_string keys). Trees are built by the unmodified reference
(or the C restatement where its library is absent), flattened by from_tree with STAGE_KEY_VARLEN and read through the
ordinary probe / scan entry points with keys packed into slots (stage_gpu_key_pack); every answer must equal
BTree::Read / RangeScanBySize on the same tree, byte for byte.
"""
import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from oracle import pyoracle

pytestmark = pytest.mark.gpu

pyoracle.build_oracle()
Table = pyoracle.RefTable if pyoracle.have_ref() else pyoracle.OracleTable
P = 8


def _random_keys(rng, n, lo, hi):
    alphabet = np.array([0x00, 0x01, 0x7F, 0x80, 0x81, 0xFF, ord("a"), ord("b")], np.uint8)
    out = set()
    while len(out) < n:
        ln = int(rng.integers(lo, hi + 1))
        out.add(bytes(alphabet[rng.integers(0, len(alphabet), ln)]))
    return sorted(out)


def _build(keys, seed):
    rng = np.random.default_rng(seed)
    order = rng.permutation(len(keys))
    pay = np.arange(len(keys), dtype=np.uint64).view(np.uint8).reshape(-1, P)
    t = Table(P, split_threshold=3072, merge_threshold=1024, leaf_node_size=4096)
    assert t.insert_varlen([keys[i] for i in order], pay[order], 0) == len(keys)
    return t


def _check_reads(t, snap, slot, keys, extra_probes, extra_starts, seed):
    rng = np.random.default_rng(seed)
    probes = [keys[i] for i in rng.integers(0, len(keys), 4000)] + extra_probes
    found, ts, kout, pay = snap.probe(sg.pack_keys(probes, slot), 0)
    wf, wts, wpay = t.probe_varlen(probes, 0)
    assert np.array_equal(found, wf) and np.array_equal(ts[wf == 1], wts[wf == 1])
    assert np.array_equal(pay[wf == 1], wpay[wf == 1])
    hit = np.flatnonzero(wf == 1)
    assert sg.unpack_keys(kout[hit]) == [probes[i] for i in hit]
    assert found[:4000].all()
    starts = [keys[i] for i in rng.integers(0, len(keys), 300)] + extra_starts
    for size in (1, 10, 100, 1000):
        counts, kout, tso, pay = snap.scan(sg.pack_keys(starts, slot), size, 0, 0)
        wc, wkeys, wts, wpay = t.scan_varlen(starts, size, 0, 0)
        assert np.array_equal(counts, wc), size
        for i in range(len(starts)):
            c = int(wc[i])
            assert sg.unpack_keys(kout[i, :c]) == wkeys[i], (size, starts[i])
            assert np.array_equal(pay[i, :c], wpay[i, :c]) and np.array_equal(tso[i, :c], wts[i, :c])


def test_to_string_keys_like_the_reference_fixture():
    """testing_btree.cpp Insert (:393-440: 100 000 std::to_string keys, full read-back), Read (:442-470: miss / hit) and
    RangeScanBySize (:630-675: 100 and 1000 rows from "9000")."""
    n = 100_000
    keys = [str(i).encode() for i in range(n)]
    t = _build(keys, 1)
    snap = sg.Snapshot.from_tree(t.root, 8, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN)       # keys of <= 7 bytes
    found, _, kout, pay = snap.probe(sg.pack_keys(keys, 8), 0)
    assert found.all() and np.array_equal(pay.view(np.uint64).ravel(), np.arange(n, dtype=np.uint64))
    assert sg.unpack_keys(kout[:2000]) == keys[:2000]
    _check_reads(t, snap, 8, keys, [b"100000", b"", b"abc", b"99", b"1000000", b"10\x80"], [b"", b"9000", b"99999", b":", b"1"], 2)
    for size in (100, 1000):
        counts, kout, _, _ = snap.scan(sg.pack_keys([b"9000"], 8), size, 0, 0)
        assert counts[0] == size                                          # the reference test pins the count
        got = sg.unpack_keys(kout[0, :size])
        assert got[0] == b"9000" and got == sorted(got) and len(set(got)) == size
    snap.close()
    t.close()


def test_byte_keys_with_prefixes_and_minimum_bytes():
    """keys of 1..15 bytes over an alphabet with 0x80 (the minimum of the signed order), 0x00, 0xFF: proper prefixes and
    keys that differ only by trailing minimum bytes are everywhere; slot of 16 bytes (the fused short-scan path)."""
    rng = np.random.default_rng(5)
    keys = _random_keys(rng, 30000, 1, 15)
    t = _build(keys, 3)
    snap = sg.Snapshot.from_tree(t.root, 16, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN)
    assert snap.info().n_rows == len(keys)
    _check_reads(t, snap, 16, keys, [b"", b"\x80", keys[7] + b"\x80", keys[9][:-1] + b"\x7f", b"zzzz"],
                 [b"", b"\x80", b"\x80\x80", b"\x7f" * 15, b"\xff", keys[11] + b"\x80"], 4)
    # device buffers and a snapshot file round trip keep the key form
    import os
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "varlen.snap")
        snap.save(path)
        again = sg.Snapshot.load(path)
        _check_reads(t, again, 16, keys, [b""], [b""], 6)
        again.close()
    snap.close()
    t.close()


def test_long_keys_in_memcmp_order():
    """keys of 16..31 bytes: KeyCompare uses memcmp for every pair; slot of 32 bytes, padding 0x00."""
    rng = np.random.default_rng(8)
    keys = _random_keys(rng, 20000, 16, 31)
    t = _build(keys, 9)
    snap = sg.Snapshot.from_tree(t.root, 32, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN)
    _check_reads(t, snap, 32, keys, [keys[3] + b"\x00", keys[4][:-1], b"\xff" * 31], [b"\x00" * 16, b"\xff" * 20, keys[5] + b"\x00"], 10)
    snap.close()
    t.close()


def test_tables_without_a_single_order_are_rejected():
    keys = [b"short", b"a-key-of-sixteen!", b"another-key-of-more-than-16"]
    t = _build(keys, 11)
    with pytest.raises(sg.StageError):                   # lengths straddle 16: KeyCompare has no single order
        sg.Snapshot.from_tree(t.root, 32, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN)
    with pytest.raises(sg.StageError):                   # signed tables hold keys of <= 15 bytes
        sg.Snapshot.from_tree(t.root, 32, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN | sg.KEY_SIGNED_BYTES)
    t.close()
    t = _build([str(i).encode() for i in range(5000)], 12)
    with pytest.raises(sg.StageError):                   # slot too small for "1000".."4999"
        sg.Snapshot.from_tree(t.root, 4, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN)
    with pytest.raises(sg.StageError):                   # fixed-length schema on a mixed-length tree
        sg.Snapshot.from_tree(t.root, 4, P, leaf_node_size=4096)
    t.close()


def test_refresh_of_a_variable_length_table():
    """Structural refresh (two-table and in-place) on a table with keys of different lengths: the walker packs the new
    separators, the transform the new records; answers equal the tree's and a from-scratch snapshot's."""
    keys = [str(i).encode() for i in range(0, 60000, 2)]
    t = _build(keys, 21)
    sg.set_row_slack(0.6)
    try:
        snap = sg.Snapshot.from_tree(t.root, 8, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN)
    finally:
        sg.set_row_slack(0.0)
    try:
        more = [str(i).encode() for i in range(1, 20001, 2)]                  # odd numbers: inserts everywhere, splits
        pay = (np.arange(len(more), dtype=np.uint64) * 2 + 1).view(np.uint8).reshape(-1, P)
        assert t.insert_varlen(more, pay, 0) == len(more)
        all_keys = keys + more
        s2, st = snap.refresh_from_tree(t.root)
        assert st.rows == len(all_keys) and st.leaves_transformed > 0
        _check_reads(t, s2, 8, all_keys, [b"", b"60001", b"x"], [b"", b"1", b"19999", b"5"], 22)
        s2.close()
        st = snap.refresh_in_place(t.root)
        assert st.rows == len(all_keys)
        _check_reads(t, snap, 8, all_keys, [b"", b"60001", b"x"], [b"", b"1", b"19999", b"5"], 23)
        for k in more[:3000]:
            t.delete(np.frombuffer(k, np.uint8))
        st = snap.refresh_in_place(t.root)
        assert st.rows == len(all_keys) - 3000
        _check_reads(t, snap, 8, keys + more[3000:], [more[0], more[2999], b""], [b"", more[10], b"2"], 24)
        with sg.Snapshot.from_tree(t.root, 8, P, leaf_node_size=4096, key_kind=sg.KEY_VARLEN) as fresh:
            slots = sg.pack_keys(all_keys[::7], 8)
            a, b = snap.probe(slots, 9), fresh.probe(slots, 9)
            assert all(np.array_equal(x, y) for x, y in zip(a, b))
    finally:
        snap.close()
        t.close()
