"""ctypes wrappers over the two CPU checkers — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

* ``OracleTable``  -> oracle/_build/libstage_oracle.so  (plain-C restatement, oracle/stage_oracle.c)
* ``RefTable``     -> oracle/_ref/libstage_ref.so       (the UNMODIFIED reference compiled by
                      oracle/Makefile.ref + our harness oracle/ref_harness.cpp)

Both expose the same interface so tests can run either as the checker.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "_build", "libstage_oracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libstage_ref.so")

_vp, _u32, _u64, _i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int


def build_oracle():
    """Compile the C restatement (and the reference itself when /root/reference is present)."""
    subprocess.run(["make", "-s", "-C", _HERE, "_build/libstage_oracle.so"], check=True)
    if os.path.isdir("/root/reference/src/vstore"):
        subprocess.run(["make", "-s", "-C", _HERE, "-f", "Makefile.ref", "-j8", "all"], check=True)


def have_ref():
    return os.path.exists(REF_SO)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


class _Table:
    """Common probe/scan plumbing over a C library exposing the <prefix>_* entry points."""

    _lib = None
    prefix = ""

    def _fn(self, name, restype, argtypes):
        f = getattr(self._lib, self.prefix + name)
        f.restype, f.argtypes = restype, argtypes
        return f

    def _bind(self):
        self._insert = self._fn("insert_batch" if self.prefix == "so_" else "insert", C.c_int64,
                                [_vp, _vp, _u32, _vp, _u64, _u32])
        self._insert_varlen = self._fn("insert_varlen", C.c_int64, [_vp, _vp, _vp, _vp, _u64, _u32])
        self._load_int = self._fn("load_ycsb_int", C.c_int64, [_vp, _u64, _u64])
        self._delete = self._fn("delete", _i32, [_vp, _vp, _u32])
        self._commit_update = self._fn("commit_update", _i32, [_vp, _vp, _u32, _vp, _u32])
        self._commit_update_batch = self._fn("commit_update_batch", C.c_int64, [_vp, _vp, _u32, _u64, _u32])
        self._root = self._fn("root", _vp, [_vp])
        self._probe = self._fn("probe", _i32, [_vp, _vp, _u32, _u64, _u32, _vp, _vp, _vp, _vp])
        self._scan = self._fn("scan", _i32, [_vp, _vp, _u32, _vp, _u64, _u32, _u32, _i32, _vp, _vp, _vp, _vp])
        self._probe_varlen = self._fn("probe_varlen", _i32, [_vp, _vp, _vp, _u64, _u32, _vp, _vp, _vp])
        self._scan_varlen = self._fn("scan_varlen", _i32, [_vp, _vp, _vp, _vp, _u64, _u32, _u32, _i32, _vp, _vp, _u32, _vp, _vp, _vp])
        self._time_probe = self._fn("time_probe", C.c_double, [_vp, _vp, _u32, _u64, _u32, _i32, _i32, _vp])
        self._time_scan = self._fn("time_scan", C.c_double,
                                   [_vp, _vp, _u32, _u32, _u64, _u32, _i32, _i32, _i32, _vp])

    # -- construction ------------------------------------------------------------------------
    def insert(self, keys, payloads, commit_id=0):
        keys = np.ascontiguousarray(keys, np.uint8)
        payloads = np.ascontiguousarray(payloads, np.uint8)
        n, ks = keys.shape
        assert payloads.shape == (n, self.payload_size)
        return self._insert(self.h, _ptr(keys), ks, _ptr(payloads), n, commit_id)

    def insert_varlen(self, key_list, payloads, commit_id=0):
        blob = np.frombuffer(b"".join(key_list), np.uint8).copy()
        off = np.zeros(len(key_list) + 1, np.uint32)
        off[1:] = np.cumsum([len(k) for k in key_list])
        payloads = np.ascontiguousarray(payloads, np.uint8)
        return self._insert_varlen(self.h, _ptr(blob), _ptr(off), _ptr(payloads), len(key_list), commit_id)

    def load_ycsb_int(self, begin, end):
        return self._load_int(self.h, begin, end)

    def delete(self, key):
        key = np.ascontiguousarray(key, np.uint8)
        return self._delete(self.h, _ptr(key), key.size)

    def commit_update(self, key, new_payload, commit_id):
        key = np.ascontiguousarray(key, np.uint8)
        new_payload = np.ascontiguousarray(new_payload, np.uint8)
        assert new_payload.size == self.payload_size
        return self._commit_update(self.h, _ptr(key), key.size, _ptr(new_payload), commit_id)

    def commit_update_batch(self, keys, first_commit_id):
        """n committed YCSB-style updates (column 1 <- (uint8) commit id), commit ids first..first+n-1."""
        keys = np.ascontiguousarray(keys, np.uint8)
        n, ks = keys.shape
        return self._commit_update_batch(self.h, _ptr(keys), ks, n, first_commit_id)

    @property
    def root(self):
        return self._root(self.h)

    # -- reads -------------------------------------------------------------------------------
    def probe(self, keys, read_ts):
        keys = np.ascontiguousarray(keys, np.uint8)
        n, ks = keys.shape
        found = np.zeros(n, np.uint8)
        ts = np.zeros(n, np.uint32)
        key_out = np.zeros((n, ks), np.uint8)
        pay = np.zeros((n, self.payload_size), np.uint8)
        self._probe(self.h, _ptr(keys), ks, n, read_ts, _ptr(found), _ptr(ts), _ptr(key_out), _ptr(pay))
        return found, ts, key_out, pay

    def scan(self, start_keys, scan_sizes, read_ts, apply_visibility, want_payload=True):
        start_keys = np.ascontiguousarray(start_keys, np.uint8)
        n, ks = start_keys.shape
        scan_sizes = np.ascontiguousarray(np.broadcast_to(np.asarray(scan_sizes, np.uint32), (n,)))
        stride = int(scan_sizes.max()) if n else 0
        counts = np.zeros(n, np.uint32)
        key_out = np.zeros((n, stride, ks), np.uint8)
        ts = np.zeros((n, stride), np.uint32)
        pay = np.zeros((n, stride, self.payload_size), np.uint8) if want_payload else None
        self._scan(self.h, _ptr(start_keys), ks, _ptr(scan_sizes), n, stride, read_ts, int(apply_visibility),
                   _ptr(counts), _ptr(key_out), _ptr(ts), _ptr(pay))
        return counts, key_out, ts, pay

    # -- the same reads for keys of different lengths (lists of bytes) --------------------------
    @staticmethod
    def _blob(key_list):
        blob = np.frombuffer(b"".join(key_list) + b"\0", np.uint8).copy()
        off = np.zeros(len(key_list) + 1, np.uint32)
        off[1:] = np.cumsum([len(k) for k in key_list])
        return blob, off

    def probe_varlen(self, key_list, read_ts):
        """-> (found u8[n], ts u32[n], payload u8[n, P])"""
        blob, off = self._blob(key_list)
        n = len(key_list)
        found, ts = np.zeros(n, np.uint8), np.zeros(n, np.uint32)
        pay = np.zeros((n, self.payload_size), np.uint8)
        self._probe_varlen(self.h, _ptr(blob), _ptr(off), n, read_ts, _ptr(found), _ptr(ts), _ptr(pay))
        return found, ts, pay

    def scan_varlen(self, start_list, scan_sizes, read_ts, apply_visibility, key_stride=32):
        """-> (row_counts u32[n], keys: list of lists of bytes, ts u32[n, S], payload u8[n, S, P])"""
        blob, off = self._blob(start_list)
        n = len(start_list)
        scan_sizes = np.ascontiguousarray(np.broadcast_to(np.asarray(scan_sizes, np.uint32), (n,)))
        stride = int(scan_sizes.max()) if n else 0
        counts = np.zeros(n, np.uint32)
        key_out = np.zeros((n, stride, key_stride), np.uint8)
        lens = np.zeros((n, stride), np.uint32)
        ts = np.zeros((n, stride), np.uint32)
        pay = np.zeros((n, stride, self.payload_size), np.uint8)
        self._scan_varlen(self.h, _ptr(blob), _ptr(off), _ptr(scan_sizes), n, stride, read_ts, int(apply_visibility),
                          _ptr(counts), _ptr(key_out), key_stride, _ptr(lens), _ptr(ts), _ptr(pay))
        keys = [[bytes(key_out[i, j, : lens[i, j]]) for j in range(counts[i])] for i in range(n)]
        return counts, keys, ts, pay

    def time_probe(self, keys, read_ts, threads, pin=True):
        keys = np.ascontiguousarray(keys, np.uint8)
        n, ks = keys.shape
        hits = C.c_uint64(0)
        sec = self._time_probe(self.h, _ptr(keys), ks, n, read_ts, threads, int(pin), C.byref(hits))
        return sec, hits.value

    def time_scan(self, start_keys, scan_size, read_ts, apply_visibility, threads, pin=True):
        start_keys = np.ascontiguousarray(start_keys, np.uint8)
        n, ks = start_keys.shape
        rows = C.c_uint64(0)
        sec = self._time_scan(self.h, _ptr(start_keys), ks, scan_size, n, read_ts, int(apply_visibility), threads,
                              int(pin), C.byref(rows))
        return sec, rows.value


class OracleTable(_Table):
    prefix = "so_"

    def __init__(self, payload_size, split_threshold=16384, merge_threshold=32768, leaf_node_size=65536):
        if OracleTable._lib is None:
            if not os.path.exists(ORACLE_SO):
                build_oracle()
            OracleTable._lib = C.CDLL(ORACLE_SO)
        self.payload_size = payload_size
        self.leaf_node_size = leaf_node_size
        self._bind()
        create = self._fn("tree_create", _vp, [_u32, _u32, _u32, _u32])
        self._destroy = self._fn("tree_destroy", None, [_vp])
        self._walk = self._fn("walk", _u64, [_vp, _vp, _vp, _vp, _u64])
        self.h = create(payload_size, split_threshold, merge_threshold, leaf_node_size)

    def close(self):
        if self.h:
            self._destroy(self.h)
            self.h = None

    def walk(self, root=None):
        """In-order leaf pointers and separators of any tree in the reference layout."""
        return walk_tree(self.root if root is None else root)


class RefTable(_Table):
    prefix = "ref_"

    def __init__(self, payload_size, split_threshold=16384, merge_threshold=32768, leaf_node_size=65536,
                 catalog_key_len=8, leaf_pool_blocks=1 << 20):
        if RefTable._lib is None:
            RefTable._lib = C.CDLL(REF_SO)
        self.payload_size = payload_size
        self.leaf_node_size = leaf_node_size
        self._bind()
        create = self._fn("table_create", _vp, [_u32, _u32, _u32, _u32, _u32, _u64])
        self.h = create(catalog_key_len, payload_size, split_threshold, merge_threshold, leaf_node_size,
                        leaf_pool_blocks)

    def close(self):
        self.h = None  # the reference has no working table destructor path; pools are leaked

    # -- in-flight updates (live reference only) ------------------------------------
    def begin_update(self, key, new_payload, txn_id):
        """LeafNode::Update without the finalize: control bit set, undo copy in the EphemeralPool, payload
        overwritten in place. Returns 0 on success (the reference's ReturnCode byte otherwise)."""
        key = np.ascontiguousarray(key, np.uint8)
        new_payload = np.ascontiguousarray(new_payload, np.uint8)
        assert new_payload.size == self.payload_size
        f = self._fn("begin_update", _i32, [_vp, _vp, _u32, _vp, _u32])
        return f(self.h, _ptr(key), key.size, _ptr(new_payload), txn_id)

    def undo_resolver(self):
        """(function address, context) of the stage_undo_resolver_t over this table's EphemeralPool."""
        return C.cast(getattr(self._lib, "ref_resolve_undo"), _vp).value, int(self.h)


class Writer:
    """A thread running real update transactions on a RefTable (ref_update_loop: LeafNode::Update + the commit install,
    in the reference's order) while the caller takes snapshots. ctypes releases the GIL, so it truly runs beside the
    caller. stop() -> (tried, committed u8[n])."""

    def __init__(self, table, keys_u32, first_cid):
        import threading
        self.keys = np.ascontiguousarray(keys_u32, np.uint32)
        self.first_cid = first_cid
        self.committed = np.zeros(len(self.keys), np.uint8)
        self._stop = C.c_int(0)
        self.tried = 0
        f = table._fn("update_loop", C.c_uint64, [_vp, _vp, _u32, _u64, _u32, _vp, _vp])

        def run():
            self.tried = f(table.h, _ptr(self.keys), 4, len(self.keys), first_cid, C.byref(self._stop), _ptr(self.committed))
        self._t = threading.Thread(target=run)
        self._t.start()

    def alive(self):
        return self._t.is_alive()

    def stop(self):
        self._stop.value = 1
        self._t.join()
        return int(self.tried), self.committed


def key_compare(a, b):
    """BaseNode::KeyCompare (b_tree.h:116-134) on two byte strings, as restated by the C oracle -> -1 / 0 / 1."""
    if OracleTable._lib is None:
        build_oracle()
        OracleTable._lib = C.CDLL(ORACLE_SO)
    f = OracleTable._lib.so_key_compare
    f.restype, f.argtypes = C.c_int, [C.c_char_p, _u32, C.c_char_p, _u32]
    r = f(bytes(a), len(a), bytes(b), len(b))
    return (r > 0) - (r < 0)


def walk_tree(root):
    """(leaf_ptrs[L] u64, sep_ptrs[L-1] u64, sep_lens[L-1] u32) of a tree in the reference node layout."""
    if OracleTable._lib is None:
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        OracleTable._lib = C.CDLL(ORACLE_SO)
    w = OracleTable._lib.so_walk
    w.restype, w.argtypes = _u64, [_vp, _vp, _vp, _vp, _u64]
    n = w(root, None, None, None, 0)
    leaves = np.zeros(n, np.uint64)
    seps = np.zeros(max(n, 1), np.uint64)
    lens = np.zeros(max(n, 1), np.uint32)
    w(root, _ptr(leaves), _ptr(seps), _ptr(lens), n)
    return leaves, seps[: n - 1], lens[: n - 1]


def read_mem(addr, nbytes):
    return np.frombuffer(C.string_at(int(addr), int(nbytes)), np.uint8)


# --------------------------------------------------------------------------------------------------
# key-stream generators restated from benchmark/benchmark_common.h (FastRandom :12-64, Zipf :67-105)
# --------------------------------------------------------------------------------------------------
class FastRandom:
    M48 = (1 << 48) - 1

    def __init__(self, seed):
        self.seed = (seed ^ 0x5DEECE66D) & self.M48

    def _next(self, bits):
        self.seed = (self.seed * 0x5DEECE66D + 0xB) & self.M48
        return self.seed >> (48 - bits)

    def next(self):
        return ((self._next(32) << 32) + self._next(32)) & 0xFFFFFFFFFFFFFFFF

    def next_u32(self):
        return self._next(32)

    def next_uniform(self):
        return ((self._next(26) << 27) + self._next(27)) / float(1 << 53)

    def next_string(self, n):
        return bytes(self._next(8) % 256 for _ in range(n))


_LCG_A, _LCG_C, _LCG_M = 0x5DEECE66D, 0xB, (1 << 48) - 1
_LCG_BLOCK = 1 << 16
_lcg_tab = None


def _lcg_table():
    """A_k, C_k with state_{i+k} = A_k * state_i + C_k (mod 2^48) for k = 1.._LCG_BLOCK."""
    global _lcg_tab
    if _lcg_tab is None:
        A = np.empty(_LCG_BLOCK, np.uint64)
        Cc = np.empty(_LCG_BLOCK, np.uint64)
        aa, cc = 1, 0
        for i in range(_LCG_BLOCK):
            aa, cc = (aa * _LCG_A) & _LCG_M, (cc * _LCG_A + _LCG_C) & _LCG_M
            A[i], Cc[i] = aa, cc
        _lcg_tab = (A, Cc)
    return _lcg_tab


def lcg_states(seed, count):
    """The first `count` 48-bit states of FastRandom(seed) (benchmark_common.h:55-62), vectorised.
    uint64 wrap-around multiplication keeps the low 48 bits exact."""
    A, Cc = _lcg_table()
    s = np.uint64((seed ^ 0x5DEECE66D) & _LCG_M)
    out = np.empty(count, np.uint64)
    done = 0
    mask = np.uint64(_LCG_M)
    with np.errstate(over="ignore"):
        while done < count:
            nb = min(_LCG_BLOCK, count - done)
            st = (A[:nb] * s + Cc[:nb]) & mask
            out[done:done + nb] = st
            s = st[-1]
            done += nb
    return out


def fastrandom_stream(seed, count, bits=32):
    """FastRandom(seed).next(bits) called `count` times."""
    return lcg_states(seed, count) >> np.uint64(48 - bits)


def uniform_keys(n_rows, count, seed=1):
    """Reference key pick rand() % (N+1) clamped to 1..N, driven by FastRandom(seed)."""
    x = fastrandom_stream(seed, count, 32)
    return (1 + x % np.uint64(n_rows)).astype(np.uint32)


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
