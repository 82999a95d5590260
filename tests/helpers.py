"""Shared table builders / comparison helpers for the tests (checker side only)."""
import numpy as np

from oracle import pyoracle


def u32_keys(a):
    return np.ascontiguousarray(np.asarray(a, np.uint32)).view(np.uint8).reshape(-1, 4)


def make_table(cls, payload_size, **kw):
    return cls(payload_size, **kw)


def ycsb_int_table(cls, n_rows, payload_size=1000, **kw):
    t = cls(payload_size, **kw)
    assert t.load_ycsb_int(0, n_rows) == n_rows
    return t


def random_table(cls, n_rows, key_size, payload_size, seed, **kw):
    """Rows with random fixed-length keys inserted in random order; payload derived from the key."""
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 256, size=(n_rows, key_size), dtype=np.uint8)
    keys = np.unique(keys, axis=0)
    rng.shuffle(keys)
    pay = np.empty((keys.shape[0], payload_size), np.uint8)
    pay[:] = (keys.astype(np.uint32).sum(axis=1) % 251).astype(np.uint8)[:, None]
    pay[:, : min(key_size, payload_size)] = keys[:, : min(key_size, payload_size)]
    t = cls(payload_size, **kw)
    assert t.insert(keys, pay, 0) == keys.shape[0]
    return t, keys, pay


def assert_probe_equal(got, want, what=""):
    names = ("found", "ts", "key", "payload")
    for g, w, nm in zip(got, want, names):
        if g is None or w is None:
            continue
        if not np.array_equal(g, w):
            bad = np.argwhere(np.atleast_1d((g != w).reshape(g.shape[0], -1).any(axis=1)))[:5].ravel()
            raise AssertionError(f"{what}: probe {nm} differs at {bad.tolist()}")


def assert_scan_equal(got, want, what=""):
    names = ("row_counts", "key", "ts", "payload")
    for g, w, nm in zip(got, want, names):
        if g is None or w is None:
            continue
        if not np.array_equal(g, w):
            bad = np.argwhere((g != w).reshape(g.shape[0], -1).any(axis=1))[:5].ravel()
            raise AssertionError(f"{what}: scan {nm} differs at scans {bad.tolist()}")
