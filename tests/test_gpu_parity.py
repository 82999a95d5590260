"""GPU parity tests: the CUDA path (through the C ABI) against the CPU checker on identical tables and
key streams. Bit-exact: found flags, commit ids, key bytes, payload bytes, scan row order and counts."""
import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from helpers import assert_probe_equal, assert_scan_equal, random_table, u32_keys, ycsb_int_table
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def checker_classes():
    cls = [pyoracle.OracleTable]
    if pyoracle.have_ref():
        cls.append(pyoracle.RefTable)      # the unmodified reference, when its prebuilt library travelled
    return cls


@pytest.fixture(scope="module", params=checker_classes(), ids=lambda c: c.__name__)
def ycsb10k(request):
    t = ycsb_int_table(request.param, 10000)
    snap = sg.Snapshot.from_tree(t.root, 4, 1000)
    yield t, snap
    snap.close()


def test_snapshot_shape(ycsb10k):
    t, snap = ycsb10k
    info = snap.info()
    assert info.n_rows == 10000
    leaves, _, _ = pyoracle.walk_tree(t.root)
    assert info.n_leaves == len(leaves)
    assert info.n_versions == 0


def test_probe_config1_uniform(ycsb10k):
    """BASELINE config #1 key stream: uniform 1..N (key N misses), read_ts = 1."""
    t, snap = ycsb10k
    keys = u32_keys(pyoracle.uniform_keys(10000, 1 << 16, seed=1))
    assert_probe_equal(snap.probe(keys, 1), t.probe(keys, 1), "uniform")


def test_probe_zipf_and_misses(ycsb10k):
    t, snap = ycsb10k
    z = pyoracle.Zipf(9999, 0.2, seed=1).draw(20000).astype(np.uint32)
    miss = np.array([10000, 10001, 0xFFFFFFFF, 0x80808080, 70000], np.uint32)
    keys = u32_keys(np.concatenate([z, miss, np.arange(0, 300, dtype=np.uint32)]))
    got, want = snap.probe(keys, 1), t.probe(keys, 1)
    assert_probe_equal(got, want, "zipf+miss")
    assert got[0][20000:20005].sum() == 0


def test_probe_empty_and_ragged_batches(ycsb10k):
    t, snap = ycsb10k
    for n in (0, 1, 31, 33, 255):
        keys = u32_keys(np.arange(n, dtype=np.uint32) * 7)
        assert_probe_equal(snap.probe(keys, 0), t.probe(keys, 0), f"n={n}")


@pytest.mark.parametrize("scan_size", [1, 10, 100, 1000])
@pytest.mark.parametrize("vis", [0, 1])
def test_scan_ycsb(ycsb10k, scan_size, vis):
    t, snap = ycsb10k
    n = 256 if scan_size < 1000 else 64
    sk = u32_keys(pyoracle.uniform_keys(10000, n, seed=3))
    assert_scan_equal(snap.scan(sk, scan_size, 1, vis), t.scan(sk, scan_size, 1, vis), f"S={scan_size}")


def test_scan_edges(ycsb10k):
    """start keys below the smallest / above the largest key, absent keys, ragged sizes."""
    t, snap = ycsb10k
    sk = u32_keys(np.array([32896, 0x80808080, 0x7F7F7F7F, 9999, 10000, 98175, 0, 5, 123456], np.uint32))
    sizes = np.array([5, 1000, 7, 3, 50, 9, 1, 64, 33], np.uint32)
    assert_scan_equal(snap.scan(sk, sizes, 1, 0), t.scan(sk, sizes, 1, 0), "edges")
    assert_scan_equal(snap.scan(sk, sizes, 1, 1), t.scan(sk, sizes, 1, 1), "edges/vis")


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
@pytest.mark.parametrize("key_size,payload", [(16, 400), (32, 64), (8, 96), (16, 1000)])
def test_other_key_shapes(cls, key_size, payload):
    """memcmp-ordered 16/32-byte keys (YCSB string, TPC-C stock / order_line) and 8-byte signed keys."""
    t, keys, pay = random_table(cls, 3000, key_size, payload, seed=key_size + payload)
    with sg.Snapshot.from_tree(t.root, key_size, payload) as snap:
        rng = np.random.default_rng(5)
        probe = np.concatenate([keys[rng.integers(0, len(keys), 4000)],
                                rng.integers(0, 256, size=(500, key_size), dtype=np.uint8)])
        assert_probe_equal(snap.probe(probe, 1), t.probe(probe, 1), "probe")
        sk = probe[:200]
        for S in (10, 150):
            assert_scan_equal(snap.scan(sk, S, 1, 0), t.scan(sk, S, 1, 0), f"scan S={S}")


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
def test_mvcc_chains_and_deletes(cls):
    """Committed updates create TupleHeader chains; deletes leave leaves whose max key is not the
    separator (iterator stops early). Reads at timestamps straddling the commit ids."""
    N, P = 6000, 1000
    t = ycsb_int_table(cls, N, P)
    rng = np.random.default_rng(11)
    cid = 10
    for k in rng.integers(0, N, 1500):
        newp = np.full(P, cid % 251, np.uint8)
        assert t.commit_update(u32_keys([k])[0], newp, cid) == 0
        cid += 1
    for k in rng.choice(N, 300, replace=False):
        t.delete(u32_keys([k])[0])
    with sg.Snapshot.from_tree(t.root, 4, P) as snap:
        assert snap.info().n_versions >= 1000      # chains of deleted rows are not staged
        keys = u32_keys(rng.integers(0, N + 10, 8000))
        sk = u32_keys(rng.integers(0, N, 200))
        for read_ts in (0, 1, 9, 10, 500, 760, 1509, 1510, 4000):
            assert_probe_equal(snap.probe(keys, read_ts), t.probe(keys, read_ts), f"ts={read_ts}")
            for S in (10, 200):
                for vis in (0, 1):
                    assert_scan_equal(snap.scan(sk, S, read_ts, vis), t.scan(sk, S, read_ts, vis),
                                      f"ts={read_ts} S={S} vis={vis}")


def test_generated_ycsb_matches_row_function():
    """On-device table generation (stage_gpu_snapshot_generate_ycsb): probes must agree with a tree the
    checker loads normally (probe results do not depend on the tree shape)."""
    N = 50000
    t = ycsb_int_table(pyoracle.OracleTable, N)
    with sg.Snapshot.generate_ycsb(N, 1000, rows_per_leaf=61) as snap:
        assert snap.info().n_rows == N
        keys = u32_keys(pyoracle.uniform_keys(N, 30000, seed=9))
        assert_probe_equal(snap.probe(keys, 1), t.probe(keys, 1), "generated")
        # consolidated leaves: a scan is the next S keys in reference order
        k = np.arange(N, dtype=np.uint32)
        order = k[np.argsort(k.byteswap() ^ np.uint32(0x80808080), kind="stable")]
        pos = {int(v): i for i, v in enumerate(order)}
        sk = keys[:64]
        counts, key_out, ts, pay = snap.scan(sk, 100, 1, 1)
        for i in range(64):
            p = pos[int(sk[i].view(np.uint32)[0])] if int(sk[i].view(np.uint32)[0]) < N else None
            if p is None:
                continue
            want = order[p:p + 100]
            assert counts[i] == len(want)
            assert np.array_equal(key_out[i, :counts[i]].view(np.uint32).ravel(), want)
            assert np.array_equal(pay[i, :counts[i], 0], (want & 0xFF).astype(np.uint8))
            assert np.array_equal(pay[i, :counts[i], 999], (want & 0xFF).astype(np.uint8))


def test_fused_probe_kernel_variant():
    """The opt-in warp-specialised probe kernel (csrc/probe_fused.cuh, STAGE_PROBE_FUSED=1) must be bit-identical to
    the default two-kernel path: the probe tests of this file are re-run in a child process with the variant on."""
    import os
    import subprocess
    import sys
    if os.environ.get("STAGE_VARIANT_RUN"):
        pytest.skip("already inside a variant run")
    env = dict(os.environ, STAGE_PROBE_FUSED="1", STAGE_VARIANT_RUN="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k",
                        "probe or mvcc or other_key or generated"], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


def test_fence_search_probe_variant():
    """Point probes normally locate rows through the seal-time directory (csrc/probe.cuh, dir_lookup); with
    STAGE_PROBE_DIR=0 they descend the ordered fence levels like the scans do. Both must give the same bytes: the
    probe tests of this file, the goldens and the routed probes are re-run in a child process with the directory off."""
    import os
    import subprocess
    import sys
    if os.environ.get("STAGE_VARIANT_RUN"):
        pytest.skip("already inside a variant run")
    env = dict(os.environ, STAGE_PROBE_DIR="0", STAGE_VARIANT_RUN="1")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), os.path.join(here, "test_gpu_golden.py"),
                        os.path.join(here, "test_gpu_exchange.py"), "-x", "-q", "-m", "gpu", "-k",
                        "probe or mvcc or other_key or generated or golden"], env=env, capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


def test_two_kernel_short_scan_variant():
    """Scans of <= 32 rows over rows of <= 256 bytes normally run in one kernel that plans the scan and emits its rows
    (csrc/scan_short.cuh); STAGE_SCAN_FUSED=0 keeps the plan kernel + gather pair. Both must give the checker's bytes:
    the scan tests of this file, the fuzz and the TPC-C tests are re-run in a child process with the fused kernel off."""
    import os
    import subprocess
    import sys
    if os.environ.get("STAGE_VARIANT_RUN"):
        pytest.skip("already inside a variant run")
    env = dict(os.environ, STAGE_SCAN_FUSED="0", STAGE_VARIANT_RUN="1", STAGE_TEST_TPCC_W="4")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), os.path.join(here, "test_gpu_fuzz.py"),
                        os.path.join(here, "test_gpu_tpcc.py"), "-x", "-q", "-m", "gpu", "-k", "scan or fuzz or stock_level or other_key"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


def test_column_pushdown_unaligned(ycsb10k):
    """Projection of an unaligned byte range, including versions served from the staged chains."""
    t, snap = ycsb10k
    keys = u32_keys(pyoracle.uniform_keys(10000, 5000, seed=4))
    f, ts, col = snap.probe_column(keys, 1, 3, 7)
    wf, wts, _, wp = t.probe(keys, 1)
    assert np.array_equal(f, wf) and np.array_equal(ts, wts)
    assert np.array_equal(col[wf == 1], wp[wf == 1][:, 3:10]) and not col[wf == 0].any()
    counts, kout, scol = snap.scan_column(keys[:40], 100, 1, 0, 996, 4)
    wc, wk, _, wpay = t.scan(keys[:40], 100, 1, 0)
    assert np.array_equal(counts, wc) and np.array_equal(kout, wk)
    valid = np.arange(100)[None, :] < wc[:, None]
    assert np.array_equal(scol[valid], wpay[valid][:, 996:1000])
    with pytest.raises(sg.StageError):
        snap.probe_column(keys, 1, 998, 4)


@pytest.mark.parametrize("cls", checker_classes(), ids=lambda c: c.__name__)
def test_empty_and_tiny_tables(cls):
    """A tree with no rows (the root is an empty leaf), one row, and a tree emptied again by deletes."""
    P = 64
    probes = u32_keys(np.array([0, 1, 2, 77, 0xFFFFFFFF], np.uint32))
    t = cls(P)
    with sg.Snapshot.from_tree(t.root, 4, P) as snap:
        assert snap.info().n_rows == 0
        assert_probe_equal(snap.probe(probes, 5), t.probe(probes, 5), "empty probe")
        assert_scan_equal(snap.scan(probes, 10, 5, 1), t.scan(probes, 10, 5, 1), "empty scan")
    assert t.insert(u32_keys([77]), np.full((1, P), 9, np.uint8), 0) == 1
    with sg.Snapshot.from_tree(t.root, 4, P) as snap:
        assert snap.info().n_rows == 1
        assert_probe_equal(snap.probe(probes, 5), t.probe(probes, 5), "one-row probe")
        for vis in (0, 1):
            assert_scan_equal(snap.scan(probes, 3, 5, vis), t.scan(probes, 3, 5, vis), "one-row scan")
    t.delete(u32_keys([77])[0])
    with sg.Snapshot.from_tree(t.root, 4, P) as snap:
        assert snap.info().n_rows == 0
        assert_probe_equal(snap.probe(probes, 5), t.probe(probes, 5), "deleted probe")
        assert_scan_equal(snap.scan(probes, 3, 5, 0), t.scan(probes, 3, 5, 0), "deleted scan")


@pytest.mark.parametrize("key_size,payload,leaf", [(32, 3000, 65536), (4, 8, 4096), (16, 1, 65536), (12, 20, 8192)])
def test_extreme_shapes(key_size, payload, leaf):
    """Large rows (few per leaf), tiny rows (hundreds per leaf: the slot-order truncation rule bites), small leaves,
    a 12-byte key (signed-byte order, three key words)."""
    cls = pyoracle.OracleTable
    rng = np.random.default_rng(key_size * 7 + payload)
    keys = np.unique(rng.integers(0, 256, size=(4000, key_size), dtype=np.uint8), axis=0)
    rng.shuffle(keys)
    pay = rng.integers(0, 256, size=(len(keys), payload), dtype=np.uint8)
    t = cls(payload, split_threshold=min(16384, leaf // 2), leaf_node_size=leaf)
    assert t.insert(keys, pay, 0) == len(keys)
    for k in keys[rng.choice(len(keys), 200, replace=False)]:
        t.delete(k)
    with sg.Snapshot.from_tree(t.root, key_size, payload, leaf) as snap:
        probe = np.concatenate([keys[:1500], rng.integers(0, 256, size=(200, key_size), dtype=np.uint8)])
        assert_probe_equal(snap.probe(probe, 1), t.probe(probe, 1), "probe")
        for S in (1, 7, 64, 500):
            for vis in (0, 1):
                assert_scan_equal(snap.scan(probe[:150], S, 1, vis), t.scan(probe[:150], S, 1, vis), f"S={S} vis={vis}")


def test_concurrent_callers(ycsb10k):
    """SURVEY §8b threading: several worker threads (the drivers pin N of them, ycsb_workload.cpp:522-529) call the
    batch entry points on ONE snapshot at the same time; every caller must get exactly its own results."""
    import threading
    t, snap = ycsb10k
    n_threads, rounds = 8, 6
    keys = [u32_keys(pyoracle.uniform_keys(10000, 3000 + 37 * i, seed=50 + i)) for i in range(n_threads)]
    want_p = [t.probe(k, 1) for k in keys]
    want_s = [t.scan(k[:24], 50, 1, 1) for k in keys]
    errors = []

    def worker(i):
        try:
            for _ in range(rounds):
                assert_probe_equal(snap.probe(keys[i], 1), want_p[i], f"thread {i} probe")
                assert_scan_equal(snap.scan(keys[i][:24], 50, 1, 1), want_s[i], f"thread {i} scan")
        except Exception as e:          # noqa: BLE001 - reported to the main thread
            errors.append(repr(e))

    th = [threading.Thread(target=worker, args=(i,)) for i in range(n_threads)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errors, errors[:3]


@pytest.mark.parametrize("key_size,payload", [(4, 1000), (16, 400)])
def test_physical_scan(key_size, payload):
    """TableScanExecutor(scan_sz = -1): leaves in tree order, slots in slot order (executor.h:581-619). Expected order
    is read off the checker's leaf images: slot i of leaf j, visible slots only."""
    from stage_indexorganized_b200.csrc_layout import parse_leaf
    if key_size == 4:
        t = ycsb_int_table(pyoracle.OracleTable, 8000, payload)
    else:
        t, _, _ = random_table(pyoracle.OracleTable, 3000, key_size, payload, seed=3)
    leaves, _, _ = pyoracle.walk_tree(t.root)
    want_k, want_p = [], []
    for lp in leaves:
        k, p = parse_leaf(pyoracle.read_mem(lp, 65536), key_size, payload)
        want_k.append(k)
        want_p.append(p)
    want_k, want_p = np.concatenate(want_k), np.concatenate(want_p)
    with sg.Snapshot.from_tree(t.root, key_size, payload) as snap:
        keys, ts, pay = snap.scan_physical()
        assert np.array_equal(keys, want_k) and np.array_equal(pay, want_p) and not ts.any()
        k2, _, p2 = snap.scan_physical(1234, 777)
        assert np.array_equal(k2, want_k[1234:2011]) and np.array_equal(p2, want_p[1234:2011])
        with pytest.raises(sg.StageError):
            snap.scan_physical(len(want_k) - 5, 6)


def test_probe_rows_kernel_variant():
    """STAGE_PROBE_ROWS=1: directory lookup, MVCC and the payload rows in ONE kernel (the warp that found 32 rows copies
    them, csrc/probe_rows.cuh). Same bytes as the two-kernel path: the probe tests of this file, the goldens, the TPC-C and
    host-path tests re-run in a child process with the variant forced on."""
    import os
    import subprocess
    import sys
    if os.environ.get("STAGE_VARIANT_RUN"):
        pytest.skip("already inside a variant run")
    env = dict(os.environ, STAGE_PROBE_ROWS="1", STAGE_VARIANT_RUN="1", STAGE_TEST_TPCC_W="4")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), os.path.join(here, "test_gpu_golden.py"),
                        os.path.join(here, "test_gpu_tpcc.py"), os.path.join(here, "test_gpu_hostpath.py"), "-x", "-q", "-m", "gpu",
                        "-k", "probe or mvcc or other_key or golden or q2 or stock_level_shape or batching"], env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
