"""CPU: keys of different lengths in one table (BaseNode::KeyCompare's "prefix, then length" rule,
include/vstore/b_tree.h:116-134; the std::to_string keys of test/testing_btree.cpp).

The GPU path carries such keys as fixed slots (stage_gpu_key_pack, include/stage_gpu.h). Checked here, without a GPU:
the slot order IS KeyCompare's order (so every kernel may treat the table as a fixed-length-key table), pack / unpack
round-trip, and the C restatement's variable-length reads against the unmodified reference.
"""
import functools

import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from oracle import pyoracle

pyoracle.build_oracle()


def _random_keys(rng, n, lo, hi):
    """byte strings of lo..hi bytes drawn from a small alphabet that contains both extremes of either byte order, so
    that prefixes, trailing minimum bytes (0x80 signed, 0x00 unsigned) and sign flips are all common"""
    alphabet = np.array([0x00, 0x01, 0x7F, 0x80, 0x81, 0xFF, ord("a"), ord("b")], np.uint8)
    out = set()
    while len(out) < n:
        ln = int(rng.integers(lo, hi + 1))
        out.add(bytes(alphabet[rng.integers(0, len(alphabet), ln)]))
    return sorted(out)


def _slot_cmp_signed(a, b):
    x, y = a.view(np.int8).tolist(), b.view(np.int8).tolist()
    return (x > y) - (x < y)


def _slot_cmp_unsigned(a, b):
    x, y = a.tolist(), b.tolist()
    return (x > y) - (x < y)


@pytest.mark.parametrize("lo,hi,slot,cmp_slots", [(0, 15, 16, _slot_cmp_signed), (1, 7, 8, _slot_cmp_signed),
                                                   (16, 31, 32, _slot_cmp_unsigned), (16, 23, 24, _slot_cmp_unsigned)])
def test_slot_order_is_keycompare(lo, hi, slot, cmp_slots):
    rng = np.random.default_rng(lo * 100 + hi)
    keys = _random_keys(rng, 1500, lo, hi)
    rng.shuffle(keys)
    slots = sg.pack_keys(keys, slot)
    assert sg.unpack_keys(slots) == keys
    ref_order = sorted(range(len(keys)), key=functools.cmp_to_key(
        lambda i, j: pyoracle.key_compare(keys[i], keys[j])))
    for a, b in zip(ref_order, ref_order[1:]):           # strictly ascending as slots <=> the two orders are the same
        assert cmp_slots(slots[a], slots[b]) < 0, (keys[a], keys[b])
    for _ in range(3000):                                # and pairwise, on random pairs
        i, j = (int(x) for x in rng.integers(0, len(keys), 2))
        assert cmp_slots(slots[i], slots[j]) == pyoracle.key_compare(keys[i], keys[j]), (keys[i], keys[j])
    # distinct keys give distinct slots (keys that differ only by trailing minimum bytes included)
    assert len({bytes(s) for s in slots}) == len(keys)


def test_pack_rejects_keys_outside_the_table_order():
    with pytest.raises(sg.StageError):
        sg.pack_keys([b"x" * 16], 16)                    # longer than slot - 1
    with pytest.raises(sg.StageError):
        sg.pack_keys([b"short"], 32)                     # memcmp table: KeyCompare would use the signed order for it
    with pytest.raises(sg.StageError):
        sg.pack_keys([b"x" * 16], 20, sg.KEY_SIGNED_BYTES)   # signed tables hold keys of <= 15 bytes


@pytest.mark.skipif(not pyoracle.have_ref(), reason="reference library not built")
def test_variable_length_reads_restatement_vs_reference():
    """to_string keys (testing_btree.cpp:393-440) and random byte keys: probes and range scans of the C restatement
    against the unmodified BTree::Read / RangeScanBySize, small leaves so that scans cross many of them."""
    rng = np.random.default_rng(9)
    for keys in ([str(i).encode() for i in range(20000)], _random_keys(rng, 6000, 1, 15)):
        order = rng.permutation(len(keys))
        pay = np.arange(len(keys), dtype=np.uint64).view(np.uint8).reshape(-1, 8)
        tabs = []
        for cls in (pyoracle.OracleTable, pyoracle.RefTable):
            t = cls(8, split_threshold=3072, merge_threshold=1024, leaf_node_size=4096)
            assert t.insert_varlen([keys[i] for i in order], pay[order], 0) == len(keys)
            tabs.append(t)
        o, r = tabs
        probes = [keys[i] for i in rng.integers(0, len(keys), 3000)] + [b"nope", b"", b"\x80", keys[5] + b"\x80"]
        fo, _, po = o.probe_varlen(probes, 0)
        fr, _, pr = r.probe_varlen(probes, 0)
        assert np.array_equal(fo, fr) and np.array_equal(po[fo == 1], pr[fr == 1]) and fo[:3000].all()
        starts = [keys[i] for i in rng.integers(0, len(keys), 200)] + [b"", b"9000", b"\xff" * 3, b"a\x80"]
        for size in (1, 7, 100, 1000):
            co, ko, _, pyo = o.scan_varlen(starts, size, 0, 0)
            cr, kr, _, pyr = r.scan_varlen(starts, size, 0, 0)
            assert np.array_equal(co, cr) and ko == kr
            for i in range(len(starts)):
                assert np.array_equal(pyo[i, : co[i]], pyr[i, : cr[i]])
        o.close()
        r.close()
