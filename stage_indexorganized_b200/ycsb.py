"""Host-side helpers for the on-device YCSB table generator (csrc/generate.cuh).

``rank_to_key(N, r)`` is the numpy mirror of ``ycsb_rank_to_key``: the r-th key (0-based) of the YCSB
int key set {0..N-1} in REFERENCE key order (signed bytes, least-significant byte first —
include/vstore/b_tree.h:99-134 of the reference). Used to draw scan start keys / splitters that fall
inside a given range shard.
"""
import numpy as np

# ---- the drivers' generators (benchmark/benchmark_common.h:12-105), vectorised ------------------------------
_LCG_A, _LCG_C, _LCG_M = 0x5DEECE66D, 0xB, (1 << 48) - 1


_JUMP_BLOCK = 1 << 16
_jump_tab = None


def _jump_table():
    """(A_k, C_k), k = 1.._JUMP_BLOCK, with s_{i+k} = A_k s_i + C_k (mod 2^48): the k-fold composition of the LCG step."""
    global _jump_tab
    if _jump_tab is None:
        A = np.empty(_JUMP_BLOCK, np.uint64)
        Cc = np.empty(_JUMP_BLOCK, np.uint64)
        aa, cc = 1, 0
        for i in range(_JUMP_BLOCK):
            aa, cc = (aa * _LCG_A) & _LCG_M, (cc * _LCG_A + _LCG_C) & _LCG_M
            A[i], Cc[i] = aa, cc
        _jump_tab = (A, Cc)
    return _jump_tab


def lcg_states(seed, count):
    """States s_1..s_count of FastRandom(seed): s_0 = (seed ^ A) & M, s_{k+1} = (A s_k + C) & M
    (benchmark_common.h:52-60). Vectorised block by block with the jump table (uint64 products wrap mod 2^64, which
    keeps the low 48 bits exact)."""
    count = int(count)
    A, Cc = _jump_table()
    m = np.uint64(_LCG_M)
    s = np.uint64((int(seed) ^ _LCG_A) & _LCG_M)
    out = np.empty(count, np.uint64)
    done = 0
    with np.errstate(over="ignore"):
        while done < count:
            nb = min(_JUMP_BLOCK, count - done)
            st = (A[:nb] * s + Cc[:nb]) & m
            out[done:done + nb] = st
            s = st[-1]
            done += nb
    return out


def fastrandom_next64(seed, count):
    """FastRandom(seed).next() called `count` times: (next(32) << 32) + next(32)."""
    st = lcg_states(seed, 2 * count)
    return ((st[0::2] >> np.uint64(16)) << np.uint64(32)) + (st[1::2] >> np.uint64(16))


def uniform_keys(n_rows, count, seed=1):
    """1 + next_u32() % N: a uniform key stream over keys 1..N, the range the YCSB driver draws from
    (ycsb.cpp:99-103; key N is never loaded, so it misses)."""
    x = lcg_states(seed, count) >> np.uint64(16)
    return (np.uint64(1) + x % np.uint64(n_rows)).astype(np.uint32)


def random_strings(seed, count, length=16):
    """FastRandom(seed).next_string(length) called `count` times (next_char = next(8) % 256)."""
    return (lcg_states(seed, count * length) >> np.uint64(40)).astype(np.uint8).reshape(count, length)


class Zipf:
    """ZipfDistribution (benchmark_common.h:67-105) with a fixed FastRandom seed."""

    def __init__(self, n, theta, seed=1):
        self.n, self.theta, self.seed = n, theta, seed
        i = np.arange(1, n + 1, dtype=np.float64)
        self.zetan = float(np.sum(np.power(1.0 / i, theta)))
        self.zeta2 = float(1.0 + np.power(0.5, theta))

    def draw(self, count):
        st = lcg_states(self.seed, 2 * count)        # NextUniform = ((next(26) << 27) + next(27)) / 2^53
        h = st[0::2] >> np.uint64(22)
        l = st[1::2] >> np.uint64(21)
        u = ((h << np.uint64(27)) + l).astype(np.float64) / float(1 << 53)
        alpha = 1.0 / (1.0 - self.theta)
        eta = (1 - (2.0 / self.n) ** (1 - self.theta)) / (1 - self.zeta2 / self.zetan)
        uz = u * self.zetan
        out = 1 + (self.n * np.power(eta * u - eta + 1, alpha)).astype(np.uint64)
        out[uz < 1 + 0.5 ** self.theta] = 2
        out[uz < 1] = 1
        return out


def ycsb_update_patch(payload_size, commit_ids):
    """The committed YCSB update used by the benchmark and the tests: payload bytes [lo, hi) <- (uint8) commit id
    (one 100-byte column of the 10 x 100-byte tuple, ycsb_mixed.cpp:38-80). Returns (patch_off, patches)."""
    lo, hi = (100 if payload_size > 100 else 0), (200 if payload_size > 200 else payload_size)
    cids = np.asarray(commit_ids, np.uint32)
    return lo, np.repeat((cids & 0xFF).astype(np.uint8)[:, None], hi - lo, axis=1)



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


# ---- YCSB string-key table (-m), built on the device (BASELINE configs[2]: 100 M tuples do not fit a host tree) ----
def _lcg_jump_torch(start_idx, seed, torch):
    """State s_k of FastRandom(seed) for every k in start_idx (int64 CUDA tensor), by composing the step maps
    (A, C)^(2^j) over the bits of k. int64 products wrap mod 2^64, which keeps the low 48 bits exact."""
    a, c, m = _LCG_A, _LCG_C, _LCG_M
    mul = torch.ones_like(start_idx)
    add = torch.zeros_like(start_idx)
    j = 0
    top = int(start_idx.max().item()) if start_idx.numel() else 0
    while (1 << j) <= top:
        bit = ((start_idx >> j) & 1).bool()
        sa = a - (1 << 64) if a >= (1 << 63) else a
        mul = torch.where(bit, (mul * sa) & m, mul)
        add = torch.where(bit, (add * sa + c) & m, add)
        a, c = (a * a) & m, (a * c + c) & m
        j += 1
    s0 = (int(seed) ^ _LCG_A) & m
    return (mul * s0 + add) & m


def string_keys_torch(seed, first_row, n_rows, device, length=16):
    """Keys of rows [first_row, first_row + n_rows) of the YCSB string table: FastRandom(seed).next_string(16) per
    row in rowid order (ycsb_loader.cpp:107-117; next_char = next(8) % 256). uint8 [n_rows, length] CUDA tensor."""
    import torch
    rows = torch.arange(first_row, first_row + n_rows, dtype=torch.int64, device=device)
    s = _lcg_jump_torch(rows * length, seed, torch)
    out = torch.empty((n_rows, length), dtype=torch.uint8, device=device)
    for ch in range(length):
        s = (s * _LCG_A + _LCG_C) & _LCG_M
        out[:, ch] = (s >> 40).to(torch.uint8)
    return out


def _key16_words(keys, torch):
    """Two int64 per 16-byte key whose SIGNED order is the memcmp order of the key bytes."""
    b = keys.to(torch.int64)
    hi = torch.zeros(keys.shape[0], dtype=torch.int64, device=keys.device)
    lo = torch.zeros_like(hi)
    for i in range(8):
        hi = (hi << 8) | b[:, i]
        lo = (lo << 8) | b[:, 8 + i]
    sign = -(1 << 63)
    return hi ^ sign, lo ^ sign


def build_string_table_device(n_rows, payload_size=1000, seed=1804289383, rows_per_leaf=61, chunk=8_000_000):
    """The YCSB string table in flattened form, built on the current CUDA device: 16 random key bytes per row
    (seeded like the reference's first un-seeded rand(), 1804289383), payload payload_size x 'a', commit id 0; rows
    sorted in reference key order (memcmp for 16-byte keys) and bulk-loaded with alloc_sorted / seal_sorted.
    Returns (snapshot, keys_by_rowid uint8 [n,16] CUDA, order int64 [n] CUDA: order[r] = rowid of the r-th smallest key)."""
    import torch
    from . import Snapshot
    dev = torch.device("cuda", torch.cuda.current_device())
    keys = torch.empty((n_rows, 16), dtype=torch.uint8, device=dev)
    for lo in range(0, n_rows, chunk):
        m = min(chunk, n_rows - lo)
        keys[lo:lo + m] = string_keys_torch(seed, lo, m, dev)
    hi_w, lo_w = _key16_words(keys, torch)
    p1 = torch.sort(lo_w, stable=True).indices
    p2 = torch.sort(hi_w[p1], stable=True).indices
    order = p1[p2]
    del hi_w, lo_w, p1, p2
    snap, raw, pay, _ = Snapshot.alloc_sorted(n_rows, 16, payload_size, rows_per_leaf)
    for lo in range(0, n_rows, chunk):
        m = min(chunk, n_rows - lo)
        raw[lo:lo + m] = keys[order[lo:lo + m]]
    pay.fill_(ord("a"))
    snap.seal_sorted()
    return snap, keys, order
