"""GPU: seeded differential fuzz of the whole snapshot life cycle against the checker (C restatement and, where
built, the unmodified reference): random key / payload sizes and leaf geometry, then rounds of inserts, deletes and
committed updates on the host tree, each followed by a structural refresh (odd rounds) or a flattening from scratch
(even rounds), and by probes and scans of random sizes at random read ids in both visibility modes."""
import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from helpers import assert_probe_equal, assert_scan_equal
from oracle import pyoracle

pytestmark = pytest.mark.gpu

SHAPES = [  # key bytes, payload bytes, split threshold, leaf node size
    (4, 1000, 16384, 65536), (4, 13, 4096, 8192), (8, 95, 8192, 16384), (12, 60, 4096, 8192),
    (16, 200, 16384, 65536), (16, 400, 65536, 65536), (32, 60, 8192, 65536), (24, 33, 2048, 4096),
]


def _cases():
    out = []
    for i, shape in enumerate(SHAPES):
        out.append((i, shape, pyoracle.OracleTable))
        if pyoracle.have_ref() and i % 2 == 0:
            out.append((i, shape, pyoracle.RefTable))
    return out


def _rand_keys(rng, n, K, narrow):
    """Random fixed-length keys; `narrow` keeps most bytes constant so that neighbours share long prefixes."""
    k = rng.integers(0, 256, size=(n, K), dtype=np.uint8)
    if narrow:
        k[:, : K - 2] = rng.integers(0, 256, size=(1, K - 2), dtype=np.uint8)
    return k


@pytest.mark.parametrize("seed,shape,cls", _cases(), ids=lambda v: getattr(v, "__name__", str(v)))
def test_fuzz_lifecycle(seed, shape, cls):
    K, P, split, leaf = shape
    rng = np.random.default_rng(1000 + seed)
    t = cls(P, split_threshold=split, merge_threshold=split * 2, leaf_node_size=leaf)
    n0 = int(rng.integers(300, 5000))
    keys = np.unique(_rand_keys(rng, n0, K, narrow=seed % 3 == 0), axis=0)
    rng.shuffle(keys)
    pay = rng.integers(0, 256, size=(keys.shape[0], P), dtype=np.uint8)
    assert t.insert(keys, pay, 0) == keys.shape[0]
    live = [bytes(k) for k in keys]
    snap = sg.Snapshot.from_tree(t.root, K, P, leaf_node_size=leaf)
    cid = 5
    try:
        for rnd in range(5):
            # inserts
            more = _rand_keys(rng, int(rng.integers(0, 800)), K, narrow=seed % 3 == 0)
            if more.shape[0]:
                mpay = rng.integers(0, 256, size=(more.shape[0], P), dtype=np.uint8)
                t.insert(more, mpay, cid)
                live = list(set(live) | {bytes(k) for k in more})
                cid += int(rng.integers(1, 4))
            # deletes
            nd = int(rng.integers(0, max(1, len(live) // 8)))
            for j in sorted(rng.choice(len(live), nd, replace=False), reverse=True):
                t.delete(np.frombuffer(live[j], np.uint8))
                live.pop(j)
            # committed updates (some rows several times)
            nu = int(rng.integers(0, 400))
            for j in rng.integers(0, len(live), nu):
                assert t.commit_update(np.frombuffer(live[j], np.uint8), rng.integers(0, 256, P, dtype=np.uint8), cid) == 0
                cid += 1
            if rnd % 2 == 0:
                nxt, st = snap.refresh_from_tree(t.root)
                assert st.rows == len(live)
            else:
                nxt = sg.Snapshot.from_tree(t.root, K, P, leaf_node_size=leaf)
            snap.close()
            snap = nxt
            assert snap.info().n_rows == len(live)
            # reads
            lk = np.frombuffer(b"".join(live), np.uint8).reshape(-1, K)
            probes = np.concatenate([lk[rng.integers(0, len(live), 1500)], _rand_keys(rng, 300, K, False)])
            starts = np.concatenate([lk[rng.integers(0, len(live), 60)], _rand_keys(rng, 12, K, False),
                                     np.zeros((1, K), np.uint8), np.full((1, K), 255, np.uint8), np.full((1, K), 127, np.uint8),
                                     np.full((1, K), 128, np.uint8)])
            sizes = rng.integers(0, 300, starts.shape[0]).astype(np.uint32)
            sizes[:3] = (1, 2, 299)
            for read_ts in {0, cid, int(rng.integers(0, cid)), int(rng.integers(0, cid)), 0xFFFFFFFF}:
                what = f"shape {shape} round {rnd} ts {read_ts}"
                assert_probe_equal(snap.probe(probes, read_ts), t.probe(probes, read_ts), what)
                for vis in (0, 1):
                    assert_scan_equal(snap.scan(starts, sizes, read_ts, vis), t.scan(starts, sizes, read_ts, vis),
                                      f"{what} vis {vis}")
    finally:
        snap.close()
        t.close()
