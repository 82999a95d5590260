"""Host-side helpers for the on-device YCSB table generator (csrc/generate.cuh).

``rank_to_key(N, r)`` is the numpy mirror of ``ycsb_rank_to_key``: the r-th key (0-based) of the YCSB
int key set {0..N-1} in REFERENCE key order (signed bytes, least-significant byte first —
include/vstore/b_tree.h:99-134 of the reference). Used to draw scan start keys / splitters that fall
inside a given range shard.
"""
import numpy as np


def _le_count(t, al):
    t = t.astype(np.int64)
    al = al.astype(np.int64)
    t_hi = np.minimum(t, 128)
    c = np.where((al >= 128) & (t_hi > 0), np.minimum(al, 127 + t_hi) - 128 + 1, 0)
    c = c + np.where(t > 128, np.minimum(al, t - 129) + 1, 0)
    return c


def rank_to_key(n_total, ranks):
    r = np.asarray(ranks, dtype=np.int64).copy()
    shape = r.shape
    r = r.ravel()
    low = np.zeros_like(r)
    m = np.int64(n_total - 1)
    for d in range(4):
        shift = 8 * d
        a = (m - low) >> shift
        ah, al = a >> 8, a & 255
        lo = np.zeros_like(r)
        hi = np.full_like(r, 255)
        for _ in range(8):
            mid = (lo + hi + 1) >> 1
            before = mid * ah + _le_count(mid, al)
            ok = before <= r
            lo = np.where(ok, mid, lo)
            hi = np.where(ok, hi, mid - 1)
        r = r - (lo * ah + _le_count(lo, al))
        low = low | ((lo ^ 0x80) << shift)
    return low.astype(np.uint32).reshape(shape)


def sort_keys(keys):
    """uint32 keys sorted in reference order."""
    k = np.asarray(keys, np.uint32)
    return k[np.argsort(k.byteswap() ^ np.uint32(0x80808080), kind="stable")]
