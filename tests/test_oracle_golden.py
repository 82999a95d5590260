"""CPU tests (no GPU): pin the C oracle (oracle/stage_oracle.c) against
 (a) the committed golden fixtures, which are outputs of the unmodified reference (tests/golden/), and
 (b) the reference's own known-answer tests for the path (test/testing_btree.cpp), restated here, and
 (c) the live reference library when oracle/_ref is present (leaf images byte-for-byte, probes, scans).
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import cases  # noqa: E402
from oracle import pyoracle  # noqa: E402


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_oracle_matches_reference_golden(name):
    want = np.load(os.path.join(HERE, "golden", name + ".npz"))
    case = cases.CASES[name](pyoracle.OracleTable)
    got = cases.run_case(case, case["table"])
    assert sorted(got) == sorted(want.files)
    for k in want.files:
        assert np.array_equal(got[k], want[k]), f"{name}:{k}"
    case["table"].close()


def test_key_compare_semantics():
    """BaseNode::KeyCompare (include/vstore/b_tree.h:99-134): signed bytes below 16 B, memcmp from 16 B."""
    lib = pyoracle.OracleTable(8)._lib
    cmp_ = lib.so_key_compare
    cmp_.restype = int.__class__ and __import__("ctypes").c_int
    import ctypes as C
    cmp_.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32]

    def c(a, b):
        r = cmp_(a, len(a), b, len(b))
        return (r > 0) - (r < 0)

    le = lambda v: int(v).to_bytes(4, "little")
    assert c(le(255), le(256)) < 0          # 0xFF is -1 as signed char
    assert c(le(1), le(256)) > 0
    assert c(le(128), le(1)) < 0
    assert c(b"\x80" * 16, b"\x01" * 16) > 0   # >= 16 bytes: unsigned memcmp
    assert c(b"\x80" * 15, b"\x01" * 15) < 0   # < 16 bytes: signed
    assert c(b"10", b"100") < 0 and c(b"100", b"10") > 0 and c(b"10", b"10") == 0
    assert cmp_(None, 0, b"a", 1) < 0 and cmp_(b"a", 1, None, 0) > 0


def _payload_u64(vals):
    return np.asarray(vals, np.uint64).view(np.uint8).reshape(-1, 8)


def test_reference_btree_kats():
    """test/testing_btree.cpp restated on the oracle: Insert :393-440 (100 000 std::to_string keys, full
    read-back), Read :442-470 (miss / hit / miss), Delete :585-627, RangeScanBySize :630-675."""
    t = pyoracle.OracleTable(8, split_threshold=3072, merge_threshold=1024, leaf_node_size=4096)
    n = 100000
    keys = [str(i).encode() for i in range(n)]
    assert t.insert_varlen(keys, _payload_u64(range(n)), 0) == n
    # full read-back through the probe path, one key length at a time (probe() takes fixed-size batches)
    for ln in range(1, 6):
        ids = np.array([i for i in range(n) if len(keys[i]) == ln], np.uint64)
        kb = np.array([list(keys[i]) for i in ids], np.uint8)
        found, ts, ko, pay = t.probe(kb, 0)
        assert found.all()
        assert np.array_equal(pay.view(np.uint64).ravel(), ids)
    t.close()

    t = pyoracle.OracleTable(8, split_threshold=3072, merge_threshold=1024, leaf_node_size=4096)
    k10 = np.frombuffer(b"10", np.uint8)[None, :]
    k11 = np.frombuffer(b"11", np.uint8)[None, :]
    assert t.probe(k10, 0)[0][0] == 0                               # miss before insert
    dummy = [str(i).encode() for i in range(0, 100, 10)]
    t.insert_varlen(dummy, _payload_u64(range(0, 100, 10)), 0)
    f, _, _, pay = t.probe(k10, 0)
    assert f[0] == 1 and pay.view(np.uint64)[0, 0] == 10           # "10" -> 10
    assert t.probe(k11, 0)[0][0] == 0                               # "11" misses
    t.close()

    t = pyoracle.OracleTable(8, split_threshold=3072, merge_threshold=1024, leaf_node_size=4096)
    ks = [str(i).encode() for i in range(10, 60)]                   # 50 two-byte keys
    t.insert_varlen(ks, _payload_u64(range(10, 60)), 0)
    for i in range(10, 50):
        assert t.delete(np.frombuffer(str(i).encode(), np.uint8)) == 0
    kb = np.array([list(k) for k in ks], np.uint8)
    found = t.probe(kb, 0)[0]
    assert found[:40].sum() == 0 and found[40:].all()
    t.close()

    case = cases.case_btree_kat(pyoracle.OracleTable)
    t = case["table"]
    k9000 = np.frombuffer(b"9000", np.uint8)[None, :]
    for size in (100, 1000):
        counts, key_out, _, pay = t.scan(k9000, size, 0, 0)
        assert counts[0] == size                                    # the reference test pins the COUNT
    t.close()


needs_ref = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref (live reference) not built")


def _canon_leaf(img):
    a = img.copy()
    a[0:8] = 0                       # vptr
    a[24:32] = 0                     # next_record_slot (atomic, host-only)
    cnt = (int(a[32:40].view(np.uint64)[0]) >> 44) & 0xFFFF
    m = a[40:40 + cnt * 24].reshape(cnt, 24)
    m[:, 8:24] = 0                   # next_ptr / loc_ptr are host pointers
    return a


@needs_ref
@pytest.mark.parametrize("n_rows", [1, 62, 63, 64, 5000, 60000])
def test_loader_builds_byte_identical_leaves(n_rows):
    o, r = pyoracle.OracleTable(1000), pyoracle.RefTable(1000)
    assert o.load_ycsb_int(0, n_rows) == r.load_ycsb_int(0, n_rows) == n_rows
    lo, so, sl = pyoracle.walk_tree(o.root)
    lr, sr, slr = pyoracle.walk_tree(r.root)
    assert len(lo) == len(lr)
    for a, b in zip(lo, lr):
        assert np.array_equal(_canon_leaf(pyoracle.read_mem(a, 65536)), _canon_leaf(pyoracle.read_mem(b, 65536)))
    for a, b, la, lb in zip(so, sr, sl, slr):
        assert la == lb and pyoracle.read_mem(a, la).tobytes() == pyoracle.read_mem(b, lb).tobytes()
    o.close()


@needs_ref
def test_oracle_vs_live_reference_random_tables():
    for key_size, payload, seed in [(4, 1000, 1), (8, 96, 2), (16, 400, 3), (32, 64, 4)]:
        rng = np.random.default_rng(seed)
        keys = np.unique(rng.integers(0, 256, size=(20000, key_size), dtype=np.uint8), axis=0)
        rng.shuffle(keys)
        pay = rng.integers(0, 256, size=(len(keys), payload), dtype=np.uint8)
        o = pyoracle.OracleTable(payload, split_threshold=65536)
        r = pyoracle.RefTable(payload, split_threshold=65536, catalog_key_len=max(key_size, 8))
        assert o.insert(keys, pay, 0) == r.insert(keys, pay, 0) == len(keys)
        upd = keys[rng.integers(0, len(keys), 3000)]
        assert o.commit_update_batch(upd, 10) == r.commit_update_batch(upd, 10) == 3000
        for k in keys[rng.choice(len(keys), 500, replace=False)]:
            assert o.delete(k) == r.delete(k) == 0
        probes = np.concatenate([keys[rng.integers(0, len(keys), 5000)],
                                 rng.integers(0, 256, size=(500, key_size), dtype=np.uint8)])
        for ts in (0, 9, 1500, 5000):
            for a, b in zip(o.probe(probes, ts), r.probe(probes, ts)):
                assert np.array_equal(a, b)
            for S in (1, 10, 300):
                for vis in (0, 1):
                    for a, b in zip(o.scan(probes[:100], S, ts, vis), r.scan(probes[:100], S, ts, vis)):
                        assert np.array_equal(a, b), (key_size, ts, S, vis)
        o.close()
