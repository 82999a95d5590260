"""stage_indexorganized_b200 — Python host mirror of the C ABI in include/stage_gpu.h.

The product is ``libstage_gpu.so`` (hand-written sm_100a CUDA behind a C ABI, see csrc/). This module
only binds it with ctypes so that tests, ``bench.py`` and Python callers drive exactly the entry
points a C++/cgo caller would. There is NO CPU fallback: if the library is missing or no CUDA device
is usable the calls raise.

Reference interface mirrored (paths relative to the reference repository):
  BTree::Read                      include/vstore/b_tree.h:812-813   -> Snapshot.probe
  BTree::RangeScanBySize/GetNext   include/vstore/b_tree.h:830-941   -> Snapshot.scan
  IndexScanExecutor / TableScanExecutor  include/execute/executor.h:342-661 (visibility on/off)
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libstage_gpu.so")

IO_HOST, IO_DEVICE, IO_COMPACT = 0, 1, 2
XCHG_DEVICE_SYNC = 1
IPC_HANDLE_BYTES = 64
KEY_AUTO, KEY_SIGNED_BYTES, KEY_MEMCMP = 0, 1, 2
KEY_VARLEN = 0x100      # flag: keys of different lengths, passed as slots (pack_keys / unpack_keys)

_vp, _u32, _u64, _i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int


class StageError(RuntimeError):
    pass


class Schema(C.Structure):
    _fields_ = [("key_size", _u32), ("payload_size", _u32), ("leaf_node_size", _u32), ("key_kind", _u32)]


class Version(C.Structure):
    _fields_ = [("begin_id", _u32), ("commit_id", _u32), ("next", _u32), ("reserved", _u32)]


class SnapshotInfo(C.Structure):
    _fields_ = [("n_rows", _u64), ("n_leaves", _u64), ("n_versions", _u64), ("device_bytes", _u64),
                ("key_words", _u32), ("n_levels", _u32), ("max_leaf_rows", _u32), ("device", _u32)]


class AggI32(C.Structure):
    _fields_ = [("found", _u64), ("below", _u64), ("min_value", C.c_int32), ("max_value", C.c_int32)]


# stage_group_agg_i32_t: one record per group of consecutive probes
GROUP_AGG_I32 = np.dtype([("found", "<u4"), ("below", "<u4"), ("min_value", "<i4"), ("min_index", "<u4"),
                          ("max_value", "<i4"), ("last_below_index", "<u4")])


class Counters(C.Structure):
    _fields_ = [("slot_order_scans", _u64), ("reserved", _u64 * 7)]


class RefreshStats(C.Structure):
    _fields_ = [("leaves_total", _u64), ("leaves_carried", _u64), ("leaves_transformed", _u64), ("rows", _u64),
                ("versions_added", _u64), ("host_bytes_staged", _u64)]


# every symbol include/stage_gpu.h declares: (restype, argtypes)
SYMBOLS = {
    "stage_gpu_init": (_i32, [_i32]),
    "stage_gpu_key_pack": (_i32, [_vp, _vp, _u32, _vp]),
    "stage_gpu_key_unpack": (_i32, [_vp, _vp, _vp, _vp]),
    "stage_gpu_snapshot_from_tree": (_i32, [_vp, C.POINTER(Schema), C.POINTER(_vp)]),
    "stage_gpu_snapshot_from_tree_live": (_i32, [_vp, C.POINTER(Schema), _vp, _vp, C.POINTER(_vp)]),
    "stage_gpu_snapshot_refresh_from_tree": (_i32, [_vp, _vp, C.POINTER(_vp), C.POINTER(RefreshStats)]),
    "stage_gpu_snapshot_compact_versions": (_i32, [_vp, _vp, _vp]),
    "stage_gpu_snapshot_refresh_in_place": (_i32, [_vp, _vp, _vp]),
    "stage_gpu_set_row_slack": (_i32, [C.c_double]),
    "stage_gpu_snapshot_refresh_from_tree_live": (_i32, [_vp, _vp, _vp, _vp, C.POINTER(_vp), C.POINTER(RefreshStats)]),
    "stage_gpu_snapshot_begin": (_i32, [C.POINTER(Schema), _u64, _u64, C.POINTER(_vp)]),
    "stage_gpu_snapshot_add_versions": (_i32, [_vp, _vp, _vp, _u64]),
    "stage_gpu_snapshot_add_leaves": (_i32, [_vp, _vp, _vp, _vp, _vp, _u64]),
    "stage_gpu_snapshot_seal": (_i32, [_vp]),
    "stage_gpu_snapshot_generate_ycsb": (_i32, [C.POINTER(Schema), _u64, _u64, _u64, _u32, C.POINTER(_vp)]),
    "stage_gpu_snapshot_alloc_sorted": (_i32, [C.POINTER(Schema), _u64, _u32, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                                               C.POINTER(_vp)]),
    "stage_gpu_snapshot_seal_sorted": (_i32, [_vp]),
    "stage_gpu_snapshot_apply_updates": (_i32, [_vp, _vp, _vp, _u32, _u32, _vp, _u64, _i32, C.POINTER(_u64)]),
    "stage_gpu_snapshot_save": (_i32, [_vp, C.c_char_p]),
    "stage_gpu_snapshot_load": (_i32, [C.c_char_p, C.POINTER(_vp)]),
    "stage_gpu_snapshot_info": (_i32, [_vp, C.POINTER(SnapshotInfo)]),
    "stage_gpu_snapshot_schema": (_i32, [_vp, C.POINTER(Schema)]),
    "stage_gpu_snapshot_destroy": (_i32, [_vp]),
    "stage_gpu_snapshot_set_stream": (_i32, [_vp, _vp]),
    "stage_gpu_snapshot_get_stream": (_vp, [_vp]),
    "stage_gpu_snapshot_enable_timing": (_i32, [_vp, _i32]),
    "stage_gpu_snapshot_last_timing": (_i32, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "stage_gpu_probe_batch": (_i32, [_vp, _vp, _u64, _u32, _vp, _vp, _vp, _vp, _i32]),
    "stage_gpu_scan_batch": (_i32, [_vp, _vp, _vp, _u64, _u32, _u32, _i32, _vp, _vp, _vp, _vp, _i32]),
    "stage_gpu_probe_project": (_i32, [_vp, _vp, _u64, _u32, _vp, _vp, _u32, _u32, _vp, _i32]),
    "stage_gpu_probe_aggregate_i32": (_i32, [_vp, _vp, _u64, _u32, _u32, C.c_int32, C.POINTER(AggI32), _i32]),
    "stage_gpu_probe_aggregate_i32_noalloc": (_i32, [_vp, _vp, _u64, _u32, _u32, C.c_int32, _vp]),
    "stage_gpu_probe_group_aggregate_i32": (_i32, [_vp, _vp, _u64, _vp, _u64, _u32, _u32, C.c_int32, _vp, _i32]),
    "stage_gpu_scan_project": (_i32, [_vp, _vp, _vp, _u64, _u32, _u32, _i32, _vp, _vp, _u32, _u32, _vp, _i32]),
    "stage_gpu_tpcc_stock_level": (_i32, [_vp, _vp, _vp, _u64, _u32, _vp, _i32]),
    "stage_gpu_scan_physical": (_i32, [_vp, _u64, _u64, _vp, _vp, _vp, _i32]),
    "stage_gpu_route": (_i32, [C.POINTER(Schema), _vp, _u32, _vp, _u64, _vp]),
    "stage_gpu_route_device": (_i32, [C.POINTER(Schema), _vp, _u32, _vp, _u64, _vp, _vp]),
    "stage_gpu_exchange_create": (_i32, [_u32, _u32, _u64, C.POINTER(Schema), C.POINTER(_vp)]),
    "stage_gpu_exchange_ipc_handle": (_i32, [_vp, _vp]),
    "stage_gpu_exchange_attach_ipc": (_i32, [_vp, _vp]),
    "stage_gpu_exchange_attach_ptrs": (_i32, [_vp, _vp]),
    "stage_gpu_exchange_arena": (_vp, [_vp]),
    "stage_gpu_exchange_post": (_i32, [_vp, _vp, _vp, _u64, _i32, _vp]),
    "stage_gpu_exchange_serve": (_i32, [_vp, _vp, _u32, _i32, _vp]),
    "stage_gpu_exchange_serve_project": (_i32, [_vp, _vp, _u32, _u32, _u32, _i32, _vp]),
    "stage_gpu_exchange_project_results": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_u32)]),
    "stage_gpu_exchange_serve_aggregate_i32": (_i32, [_vp, _vp, _u32, _u32, C.c_int32, _i32, _vp]),
    "stage_gpu_exchange_aggregate_result": (_i32, [_vp, C.POINTER(AggI32), _vp]),
    "stage_gpu_exchange_wait": (_i32, [_vp, _vp]),
    "stage_gpu_exchange_results": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_u32)]),
    "stage_gpu_exchange_create_scan": (_i32, [_u32, _u32, _u64, _u32, C.POINTER(Schema), C.POINTER(_vp)]),
    "stage_gpu_exchange_post_scans": (_i32, [_vp, _vp, _vp, _vp, _u64, _i32, _vp]),
    "stage_gpu_exchange_serve_scans": (_i32, [_vp, _vp, _u32, _i32, _i32, _vp]),
    "stage_gpu_exchange_scan_results": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                                               C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32)]),
    "stage_gpu_exchange_peer_copy": (_i32, [_vp, _u32, _u64, _u64, _vp]),
    "stage_gpu_exchange_error": (_i32, [_vp, _vp, C.POINTER(_u32)]),
    "stage_gpu_exchange_enable_timing": (_i32, [_vp, _i32]),
    "stage_gpu_exchange_last_timing": (_i32, [_vp, C.POINTER(C.c_float)]),
    "stage_gpu_exchange_destroy": (_i32, [_vp]),
    "stage_gpu_batcher_create": (_i32, [_vp, _u32, _u32, C.POINTER(_vp)]),
    "stage_gpu_batcher_probe": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp, _vp]),
    "stage_gpu_batcher_scan": (_i32, [_vp, _vp, _u32, _u32, _i32, _vp, _vp, _vp, _vp]),
    "stage_gpu_batcher_stats": (_i32, [_vp, C.POINTER(_u64), C.POINTER(_u64)]),
    "stage_gpu_batcher_destroy": (_i32, [_vp]),
    "stage_gpu_snapshot_counters": (_i32, [_vp, C.POINTER(Counters), _i32]),
    "stage_gpu_host_register": (_i32, [_vp, C.c_size_t]),
    "stage_gpu_host_unregister": (_i32, [_vp]),
    "stage_gpu_host_alloc": (_vp, [C.c_size_t]),
    "stage_gpu_host_free": (None, [_vp]),
    "stage_gpu_launch_count": (_u64, []),
    "stage_gpu_strerror": (C.c_char_p, [_i32]),
    "stage_gpu_last_error": (C.c_char_p, []),
}

_lib = None


def build(verbose=False):
    """Compile csrc/ into libstage_gpu.so (nvcc, sm_100a). Raises on failure."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc")], capture_output=not verbose, text=True)
    if r.returncode != 0:
        raise StageError("building libstage_gpu.so failed:\n" + (r.stdout or "") + (r.stderr or ""))


def lib():
    """The loaded C-ABI library; raises loudly if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise StageError(f"{LIB_PATH} is missing: run __graft_entry__.build() / make -C "
                             f"{os.path.join(_HERE, 'csrc')} (there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(l, name)
            f.restype, f.argtypes = res, args
        _lib = l
    return _lib


def _check(rc):
    if rc != 0:
        l = lib()
        raise StageError(f"stage_gpu error {rc}: {l.stage_gpu_strerror(rc).decode()} "
                         f"[{l.stage_gpu_last_error().decode()}]")


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(_vp)
    if hasattr(a, "data_ptr"):          # torch tensor on the snapshot's device
        return _vp(a.data_ptr())
    return _vp(int(a))


def set_row_slack(fraction):
    """Row capacity kept free in snapshots built from host trees from now on (for refresh_in_place)."""
    _check(lib().stage_gpu_set_row_slack(float(fraction)))


def pack_keys(keys, slot_size, key_kind=KEY_AUTO):
    """Variable-length keys -> slots (stage_gpu_key_pack): list of bytes -> uint8 [n, slot_size]."""
    sc = Schema(slot_size, 1, 65536, key_kind | KEY_VARLEN)
    out = np.zeros((len(keys), slot_size), np.uint8)
    for i, k in enumerate(keys):
        kb = np.frombuffer(bytes(k), np.uint8)
        _check(lib().stage_gpu_key_pack(C.byref(sc), _ptr(kb) if len(kb) else None, len(kb), _ptr(out[i])))
    return out


def unpack_keys(slots, key_kind=KEY_AUTO):
    """Slots uint8 [n, slot_size] -> list of bytes (stage_gpu_key_unpack)."""
    slots = np.ascontiguousarray(slots, np.uint8)
    sc = Schema(slots.shape[1], 1, 65536, key_kind | KEY_VARLEN)
    out = []
    buf = np.zeros(slots.shape[1], np.uint8)
    ln = C.c_uint32(0)
    for i in range(slots.shape[0]):
        _check(lib().stage_gpu_key_unpack(C.byref(sc), _ptr(slots[i]), _ptr(buf), C.byref(ln)))
        out.append(bytes(buf[: ln.value]))
    return out


def init(device=0):
    _check(lib().stage_gpu_init(device))


def launch_count():
    return int(lib().stage_gpu_launch_count())


class PinnedArray:
    """numpy view over pinned host memory from stage_gpu_host_alloc."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = lib().stage_gpu_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise StageError("stage_gpu_host_alloc failed")
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().stage_gpu_host_free(self.ptr)
            self.ptr = None


class Snapshot:
    """GPU-resident flattened copy of one index-organized table (immutable once built)."""

    def __init__(self, handle, schema):
        self.h = handle
        self.schema = schema

    # -- construction --------------------------------------------------------------------------
    @classmethod
    def from_tree(cls, root, key_size, payload_size, leaf_node_size=65536, key_kind=KEY_AUTO):
        sc = Schema(key_size, payload_size, leaf_node_size, key_kind)
        h = _vp()
        _check(lib().stage_gpu_snapshot_from_tree(_vp(int(root)), C.byref(sc), C.byref(h)))
        return cls(h, sc)

    @classmethod
    def from_tree_live(cls, root, key_size, payload_size, resolve_fn, resolve_ctx, leaf_node_size=65536, key_kind=KEY_AUTO):
        """Snapshot of a tree with updates in flight. resolve_fn: address of a C function
        int(void *ctx, uint64_t copy_location, uint64_t *next_tuple_header, const void **old_record) that looks the undo
        copy up in the caller's EphemeralPool (include/stage_gpu.h, stage_undo_resolver_t); resolve_ctx: its context."""
        sc = Schema(key_size, payload_size, leaf_node_size, key_kind)
        h = _vp()
        _check(lib().stage_gpu_snapshot_from_tree_live(_vp(int(root)), C.byref(sc), _vp(int(resolve_fn)),
                                                       _vp(int(resolve_ctx)), C.byref(h)))
        return cls(h, sc)

    def refresh_from_tree(self, root, resolve_fn=None, resolve_ctx=0):
        """New snapshot of the same host tree after committed inserts / updates / deletes (and the splits they
        caused); unchanged leaves are carried over on the device. Returns (snapshot, RefreshStats); self stays valid.
        With resolve_fn (see from_tree_live) update transactions may be in flight during the call."""
        h, st = _vp(), RefreshStats()
        if resolve_fn:
            _check(lib().stage_gpu_snapshot_refresh_from_tree_live(self.h, _vp(int(root)), _vp(int(resolve_fn)),
                                                                   _vp(int(resolve_ctx)), C.byref(h), C.byref(st)))
        else:
            _check(lib().stage_gpu_snapshot_refresh_from_tree(self.h, _vp(int(root)), C.byref(h), C.byref(st)))
        return type(self)(h, self.schema), st

    def compact_versions(self):
        """Drop the versions no chain reaches any more (left behind by refreshes) -> (kept, dropped)."""
        kept, dropped = _u64(0), _u64(0)
        _check(lib().stage_gpu_snapshot_compact_versions(self.h, C.byref(kept), C.byref(dropped)))
        return kept.value, dropped.value

    def refresh_in_place(self, root):
        """stage_gpu_snapshot_refresh_in_place: THIS snapshot becomes the refreshed one (payload rows shifted inside
        their buffer; no second copy of the table). Returns RefreshStats. Needs row slack (set_row_slack before the
        snapshot was built) when rows were inserted; no reads may be in flight."""
        h, st = _vp(self.h.value if hasattr(self.h, "value") else self.h), RefreshStats()
        _check(lib().stage_gpu_snapshot_refresh_in_place(C.byref(h), _vp(int(root)), C.byref(st)))
        self.h = h
        return st

    @classmethod
    def generate_ycsb(cls, n_rows_total, payload_size=1000, rows_per_leaf=61, first_rank=0, n_local=None,
                      leaf_node_size=65536):
        sc = Schema(4, payload_size, leaf_node_size, KEY_AUTO)
        h = _vp()
        if n_local is None:
            n_local = n_rows_total - first_rank
        _check(lib().stage_gpu_snapshot_generate_ycsb(C.byref(sc), n_rows_total, first_rank, n_local, rows_per_leaf,
                                                      C.byref(h)))
        return cls(h, sc)

    @classmethod
    def alloc_sorted(cls, n_rows, key_size, payload_size, rows_per_leaf=61, key_kind=KEY_AUTO, leaf_node_size=65536):
        """Bulk load, step 1: returns (snapshot, raw_keys, payload, ts) — torch uint8/int32 views of the snapshot's own
        device columns, to be filled with rows in reference key order; then call seal_sorted()."""
        import torch
        sc = Schema(key_size, payload_size, leaf_node_size, key_kind)
        h, k, p, t = _vp(), _vp(), _vp(), _vp()
        _check(lib().stage_gpu_snapshot_alloc_sorted(C.byref(sc), n_rows, rows_per_leaf, C.byref(h), C.byref(k), C.byref(p),
                                                     C.byref(t)))
        self = cls(h, sc)
        dev = torch.device("cuda", torch.cuda.current_device())
        keys = torch.as_tensor(_DeviceView(k.value, (n_rows, key_size), "|u1", self), device=dev)
        pay = torch.as_tensor(_DeviceView(p.value, (n_rows, payload_size), "|u1", self), device=dev)
        ts = torch.as_tensor(_DeviceView(t.value, (n_rows,), "<i4", self), device=dev)
        return self, keys, pay, ts

    def seal_sorted(self):
        _check(lib().stage_gpu_snapshot_seal_sorted(self.h))

    @classmethod
    def from_leaves(cls, leaf_ptrs, fences, has_fence, key_size, payload_size, leaf_node_size=65536,
                    key_kind=KEY_AUTO, versions=None, version_payloads=None, chain_heads=None, rows_hint=None):
        """Streaming construction: begin -> add_versions -> add_leaves -> seal."""
        sc = Schema(key_size, payload_size, leaf_node_size, key_kind)
        h = _vp()
        n = len(leaf_ptrs)
        if rows_hint is None:
            rows_hint = n * (leaf_node_size // 24)
        _check(lib().stage_gpu_snapshot_begin(C.byref(sc), n, rows_hint, C.byref(h)))
        self = cls(h, sc)
        if versions is not None and len(versions):
            varr = (Version * len(versions))(*versions)
            vp = np.ascontiguousarray(version_payloads, np.uint8)
            _check(lib().stage_gpu_snapshot_add_versions(h, C.cast(varr, _vp), _ptr(vp), len(versions)))
        ptrs = (C.c_void_p * n)(*[int(p) for p in leaf_ptrs])
        fences = np.ascontiguousarray(fences, np.uint8)
        has_fence = np.ascontiguousarray(has_fence, np.uint8)
        heads_ptrs = None
        keep = []
        if chain_heads is not None:
            hp = []
            for ch in chain_heads:
                if ch is None:
                    hp.append(None)
                else:
                    a = np.ascontiguousarray(ch, np.uint32)
                    keep.append(a)
                    hp.append(a.ctypes.data)
            heads_ptrs = (C.c_void_p * n)(*hp)
        _check(lib().stage_gpu_snapshot_add_leaves(h, C.cast(ptrs, _vp), _ptr(fences), _ptr(has_fence),
                                                   C.cast(heads_ptrs, _vp) if heads_ptrs is not None else None, n))
        _check(lib().stage_gpu_snapshot_seal(h))
        return self

    def apply_updates(self, keys, patches, commit_ids, patch_off=0, device=False):
        """Committed single-row updates, in order (stage_gpu_snapshot_apply_updates). keys (n,K) u8, patches
        (n,patch_len) u8, commit_ids (n,) u32 — numpy arrays, or torch CUDA tensors with device=True.
        Returns the number of updates applied (absent keys are ignored)."""
        if not device:
            keys = np.ascontiguousarray(keys, np.uint8)
            patches = np.ascontiguousarray(patches, np.uint8)
            commit_ids = np.ascontiguousarray(commit_ids, np.uint32)
        n, patch_len = int(keys.shape[0]), int(patches.shape[1])
        applied = _u64(0)
        _check(lib().stage_gpu_snapshot_apply_updates(self.h, _ptr(keys), _ptr(patches), patch_off, patch_len,
                                                      _ptr(commit_ids), n, IO_DEVICE if device else IO_HOST,
                                                      C.byref(applied)))
        return int(applied.value)

    def save(self, path):
        _check(lib().stage_gpu_snapshot_save(self.h, os.fsencode(path)))

    @classmethod
    def load(cls, path):
        h = _vp()
        _check(lib().stage_gpu_snapshot_load(os.fsencode(path), C.byref(h)))
        sc = Schema(0, 0, 0, 0)
        _check(lib().stage_gpu_snapshot_schema(h, C.byref(sc)))
        return cls(h, sc)

    def info(self):
        i = SnapshotInfo()
        _check(lib().stage_gpu_snapshot_info(self.h, C.byref(i)))
        return i

    def counters(self, reset=False):
        c = Counters()
        _check(lib().stage_gpu_snapshot_counters(self.h, C.byref(c), int(reset)))
        return c

    def set_stream(self, cuda_stream):
        _check(lib().stage_gpu_snapshot_set_stream(self.h, _vp(int(cuda_stream))))

    def enable_timing(self, on=True):
        _check(lib().stage_gpu_snapshot_enable_timing(self.h, int(on)))

    def last_timing(self):
        """(plan_ms, gather_ms) of the last device-I/O batch call (CUDA events on the snapshot stream)."""
        p, g = C.c_float(0), C.c_float(0)
        _check(lib().stage_gpu_snapshot_last_timing(self.h, C.byref(p), C.byref(g)))
        return p.value, g.value

    def close(self):
        if self.h:
            _check(lib().stage_gpu_snapshot_destroy(self.h))
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- reads: host buffers (the call a user of the reference's API makes) ---------------------
    def probe(self, keys, read_ts, want_payload=True, out=None):
        """Batched BTree::Read + point visibility. keys: (n, key_size) uint8 host array.
        Returns (found u8[n], ts u32[n], key_out u8[n,K], payload u8[n,P])."""
        keys = np.ascontiguousarray(keys, np.uint8)
        n = keys.shape[0]
        K, P = self.schema.key_size, self.schema.payload_size
        assert keys.shape == (n, K)
        if out is None:
            found = np.zeros(n, np.uint8)
            ts = np.zeros(n, np.uint32)
            key_out = np.zeros((n, K), np.uint8)
            pay = np.zeros((n, P), np.uint8) if want_payload else None
        else:
            found, ts, key_out, pay = out
        _check(lib().stage_gpu_probe_batch(self.h, _ptr(keys), n, read_ts, _ptr(found), _ptr(ts), _ptr(key_out),
                                           _ptr(pay), IO_HOST))
        return found, ts, key_out, pay

    def scan_compact(self, start_keys, scan_sizes, read_ts, apply_visibility, want_payload=True):
        """The same scans with STAGE_IO_COMPACT: rows of all scans packed back to back. Returns (row_counts u32[n],
        keys u8[R,K], ts u32[R], payload u8[R,P]) with R = row_counts.sum(); scan i owns rows
        [row_counts[:i].sum(), +row_counts[i])."""
        start_keys = np.ascontiguousarray(start_keys, np.uint8)
        n = start_keys.shape[0]
        K, P = self.schema.key_size, self.schema.payload_size
        scan_sizes = np.ascontiguousarray(np.broadcast_to(np.asarray(scan_sizes, np.uint32), (n,)))
        stride = int(scan_sizes.max()) if n else 0
        counts = np.zeros(n, np.uint32)
        key_out = np.zeros((n * stride, K), np.uint8)
        ts = np.zeros(n * stride, np.uint32)
        pay = np.zeros((n * stride, P), np.uint8) if want_payload else None
        _check(lib().stage_gpu_scan_batch(self.h, _ptr(start_keys), _ptr(scan_sizes), n, stride, read_ts,
                                          int(apply_visibility), _ptr(counts), _ptr(key_out), _ptr(ts), _ptr(pay),
                                          IO_HOST | IO_COMPACT))
        r = int(counts.sum())
        return counts, key_out[:r], ts[:r], (pay[:r] if pay is not None else None)

    def scan(self, start_keys, scan_sizes, read_ts, apply_visibility, want_payload=True, out_stride=None):
        """Batched RangeScanBySize + executor loop. Returns (row_counts u32[n], key_out u8[n,S,K],
        ts u32[n,S], payload u8[n,S,P])."""
        start_keys = np.ascontiguousarray(start_keys, np.uint8)
        n = start_keys.shape[0]
        K, P = self.schema.key_size, self.schema.payload_size
        scan_sizes = np.ascontiguousarray(np.broadcast_to(np.asarray(scan_sizes, np.uint32), (n,)))
        stride = int(out_stride if out_stride is not None else (scan_sizes.max() if n else 0))
        counts = np.zeros(n, np.uint32)
        key_out = np.zeros((n, stride, K), np.uint8)
        ts = np.zeros((n, stride), np.uint32)
        pay = np.zeros((n, stride, P), np.uint8) if want_payload else None
        _check(lib().stage_gpu_scan_batch(self.h, _ptr(start_keys), _ptr(scan_sizes), n, stride, read_ts,
                                          int(apply_visibility), _ptr(counts), _ptr(key_out), _ptr(ts), _ptr(pay),
                                          IO_HOST))
        return counts, key_out, ts, pay

    def scan_physical(self, first_row=0, n=None, want_payload=True):
        """TableScanExecutor(scan_sz = -1): rows [first_row, first_row+n) of the physical (leaf, slot) order.
        -> (keys u8[n,K], ts u32[n], payload u8[n,P])."""
        if n is None:
            n = int(self.info().n_rows) - first_row
        K, P = self.schema.key_size, self.schema.payload_size
        keys, ts = np.zeros((n, K), np.uint8), np.zeros(n, np.uint32)
        pay = np.zeros((n, P), np.uint8) if want_payload else None
        _check(lib().stage_gpu_scan_physical(self.h, first_row, n, _ptr(keys), _ptr(ts), _ptr(pay), IO_HOST))
        return keys, ts, pay

    # -- reads with column push-down: only payload bytes [col_off, col_off + col_len) come back ----
    def probe_column(self, keys, read_ts, col_off, col_len):
        """-> (found u8[n], ts u32[n], col u8[n, col_len]); host arrays."""
        keys = np.ascontiguousarray(keys, np.uint8)
        n = keys.shape[0]
        found, ts = np.zeros(n, np.uint8), np.zeros(n, np.uint32)
        col = np.zeros((n, col_len), np.uint8)
        _check(lib().stage_gpu_probe_project(self.h, _ptr(keys), n, read_ts, _ptr(found), _ptr(ts), col_off, col_len,
                                             _ptr(col), IO_HOST))
        return found, ts, col

    def probe_aggregate_i32(self, keys, read_ts, col_off, threshold, n=None):
        """Probe + reduce an int32 payload column on the device -> AggI32(found, below, min_value, max_value).
        keys: host uint8 array (n, K), or a CUDA tensor / device pointer with n given."""
        out = AggI32()
        if n is None:
            keys = np.ascontiguousarray(keys, np.uint8)
            n, flags = keys.shape[0], IO_HOST
        else:
            flags = IO_DEVICE
        _check(lib().stage_gpu_probe_aggregate_i32(self.h, _ptr(keys), n, read_ts, col_off, threshold, C.byref(out), flags))
        return out

    def probe_group_aggregate_i32(self, keys, group_offsets, read_ts, col_off, threshold, n=None, out=None):
        """Probe + reduce an int32 payload column per GROUP of consecutive probes (CH-Q2*: one group per supplier) ->
        numpy array of GROUP_AGG_I32, one record per group. keys (n, K) uint8 and group_offsets (n_groups + 1) uint64
        host arrays; or, with n given, CUDA tensors (keys, offsets as int64, out as uint8 [n_groups, 24]) and the call
        is enqueued on the snapshot's stream."""
        if n is None:
            keys = np.ascontiguousarray(keys, np.uint8)
            off = np.ascontiguousarray(group_offsets, np.uint64)
            res = np.zeros(len(off) - 1, GROUP_AGG_I32)
            _check(lib().stage_gpu_probe_group_aggregate_i32(self.h, _ptr(keys), keys.shape[0], _ptr(off), len(off) - 1, read_ts,
                                                             col_off, threshold, _ptr(res), IO_HOST))
            return res
        n_groups = int(group_offsets.numel()) - 1
        _check(lib().stage_gpu_probe_group_aggregate_i32(self.h, _ptr(keys), n, _ptr(group_offsets), n_groups, read_ts, col_off,
                                                         threshold, _ptr(out), IO_DEVICE))
        return out

    def scan_column(self, start_keys, scan_sizes, read_ts, apply_visibility, col_off, col_len):
        """-> (row_counts u32[n], key_out u8[n,S,K], col u8[n,S,col_len]); host arrays."""
        start_keys = np.ascontiguousarray(start_keys, np.uint8)
        n = start_keys.shape[0]
        scan_sizes = np.ascontiguousarray(np.broadcast_to(np.asarray(scan_sizes, np.uint32), (n,)))
        stride = int(scan_sizes.max()) if n else 0
        counts = np.zeros(n, np.uint32)
        key_out = np.zeros((n, stride, self.schema.key_size), np.uint8)
        col = np.zeros((n, stride, col_len), np.uint8)
        _check(lib().stage_gpu_scan_project(self.h, _ptr(start_keys), _ptr(scan_sizes), n, stride, read_ts,
                                            int(apply_visibility), _ptr(counts), _ptr(key_out), col_off, col_len,
                                            _ptr(col), IO_HOST))
        return counts, key_out, col

    def probe_column_device(self, keys, n, read_ts, col_off, col_len, col_out, found=None, ts=None):
        _check(lib().stage_gpu_probe_project(self.h, _ptr(keys), n, read_ts, _ptr(found), _ptr(ts), col_off, col_len,
                                             _ptr(col_out), IO_DEVICE))

    # -- reads: device buffers (torch tensors or raw device pointers) ---------------------------
    def probe_device(self, keys, n, read_ts, found=None, ts=None, key_out=None, payload_out=None):
        _check(lib().stage_gpu_probe_batch(self.h, _ptr(keys), n, read_ts, _ptr(found), _ptr(ts), _ptr(key_out),
                                           _ptr(payload_out), IO_DEVICE))

    def scan_device(self, start_keys, scan_sizes, n, out_stride, read_ts, apply_visibility, row_counts,
                    key_out=None, ts_out=None, payload_out=None):
        _check(lib().stage_gpu_scan_batch(self.h, _ptr(start_keys), _ptr(scan_sizes), n, out_stride, read_ts,
                                          int(apply_visibility), _ptr(row_counts), _ptr(key_out), _ptr(ts_out),
                                          _ptr(payload_out), IO_DEVICE))


STOCK_LEVEL_REQ = np.dtype([("w_id", "<i8"), ("d_id", "<i8"), ("next_o_id", "<i8"), ("threshold", "<i4"), ("reserved", "<i4")])


def tpcc_stock_level(order_line, stock, reqs, read_ts, out=None, n=None):
    """stage_gpu_tpcc_stock_level: reqs = numpy array of STOCK_LEVEL_REQ (host) -> uint32 distinct-item counts; or, with n
    given, reqs / out are CUDA tensors (device pointers) and the call is enqueued on order_line's stream."""
    if n is None:
        reqs = np.ascontiguousarray(reqs, STOCK_LEVEL_REQ)
        out = np.zeros(len(reqs), np.uint32)
        _check(lib().stage_gpu_tpcc_stock_level(order_line.h, stock.h, _ptr(reqs), len(reqs), read_ts, _ptr(out), IO_HOST))
        return out
    _check(lib().stage_gpu_tpcc_stock_level(order_line.h, stock.h, _ptr(reqs), n, read_ts, _ptr(out), IO_DEVICE))
    return out


class Batcher:
    """Coalesces single operations of concurrent callers into batched launches (stage_gpu_batcher_*)."""

    def __init__(self, snap, max_batch=256, max_wait_us=30):
        self.snap = snap
        self.h = _vp()
        _check(lib().stage_gpu_batcher_create(snap.h, max_batch, max_wait_us, C.byref(self.h)))

    def probe(self, key, read_ts):
        """-> (found, ts, key_out u8[K], payload u8[P])"""
        K, P = self.snap.schema.key_size, self.snap.schema.payload_size
        key = np.ascontiguousarray(key, np.uint8)
        f, t = C.c_uint8(0), _u32(0)
        ko, po = np.zeros(K, np.uint8), np.zeros(P, np.uint8)
        _check(lib().stage_gpu_batcher_probe(self.h, _ptr(key), read_ts, C.byref(f), C.byref(t), _ptr(ko), _ptr(po)))
        return int(f.value), int(t.value), ko, po

    def scan(self, start_key, scan_size, read_ts, apply_visibility):
        """-> (count, keys u8[S,K], ts u32[S], payload u8[S,P])"""
        K, P = self.snap.schema.key_size, self.snap.schema.payload_size
        start_key = np.ascontiguousarray(start_key, np.uint8)
        c = _u32(0)
        ko, to, po = np.zeros((scan_size, K), np.uint8), np.zeros(scan_size, np.uint32), np.zeros((scan_size, P), np.uint8)
        _check(lib().stage_gpu_batcher_scan(self.h, _ptr(start_key), scan_size, read_ts, int(apply_visibility), C.byref(c),
                                            _ptr(ko), _ptr(to), _ptr(po)))
        return int(c.value), ko, to, po

    def stats(self):
        b, o = _u64(0), _u64(0)
        _check(lib().stage_gpu_batcher_stats(self.h, C.byref(b), C.byref(o)))
        return int(b.value), int(o.value)

    def close(self):
        if self.h:
            _check(lib().stage_gpu_batcher_destroy(self.h))
            self.h = None


def route(keys, splitters, key_size, key_kind=KEY_AUTO):
    """owner[i] = index of the range shard owning keys[i] (splitters = largest key of each shard but the last)."""
    keys = np.ascontiguousarray(keys, np.uint8)
    splitters = np.ascontiguousarray(splitters, np.uint8).reshape(-1, key_size)
    sc = Schema(key_size, 1, 65536, key_kind)
    owner = np.zeros(keys.shape[0], np.uint32)
    _check(lib().stage_gpu_route(C.byref(sc), _ptr(splitters), splitters.shape[0] + 1, _ptr(keys), keys.shape[0],
                                 _ptr(owner)))
    return owner


def route_device(keys_dev, n, splitters, owner_dev, key_size, cuda_stream=0, key_kind=KEY_AUTO):
    """Device-side routing: owner_dev[i] = shard owning keys_dev[i] (torch tensors / device pointers)."""
    splitters = np.ascontiguousarray(splitters, np.uint8).reshape(-1, key_size)
    sc = Schema(key_size, 1, 65536, key_kind)
    _check(lib().stage_gpu_route_device(C.byref(sc), _ptr(splitters), splitters.shape[0] + 1, _ptr(keys_dev), n,
                                        _ptr(owner_dev), _vp(int(cuda_stream))))


class _DeviceView:
    """Minimal __cuda_array_interface__ carrier so torch.as_tensor can wrap library-owned device memory."""

    def __init__(self, ptr, shape, typestr, owner, strides=None):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": strides}
        self._owner = owner


class ProbeExchange:
    """Routed point probes over peer memory (include/stage_gpu.h, "routed probes over peer memory"): this rank's
    arena (inboxes + result slab) and the mappings of every peer's arena. One instance per rank."""

    def __init__(self, world, rank, capacity, key_size, payload_size, key_kind=KEY_AUTO, max_scan_size=0):
        """max_scan_size > 0 makes this a scan exchange (post_scans / serve_scans / scan_results_torch) for up to
        `capacity` scans of up to max_scan_size rows per batch."""
        self.world, self.rank, self.capacity = world, rank, capacity
        self.schema = Schema(key_size, payload_size, 65536, key_kind)
        self.h = _vp()
        if max_scan_size:
            _check(lib().stage_gpu_exchange_create_scan(world, rank, capacity, max_scan_size, C.byref(self.schema),
                                                        C.byref(self.h)))
        else:
            _check(lib().stage_gpu_exchange_create(world, rank, capacity, C.byref(self.schema), C.byref(self.h)))

    def ipc_handle(self):
        buf = (C.c_uint8 * IPC_HANDLE_BYTES)()
        _check(lib().stage_gpu_exchange_ipc_handle(self.h, C.cast(buf, _vp)))
        return bytes(buf)

    def attach_ipc(self, handles):
        """handles: list of `world` 64-byte IPC handles indexed by rank."""
        blob = b"".join(handles)
        assert len(blob) == self.world * IPC_HANDLE_BYTES
        _check(lib().stage_gpu_exchange_attach_ipc(self.h, C.cast(C.c_char_p(blob), _vp)))

    def attach_ptrs(self, arenas):
        arr = (C.c_void_p * self.world)(*[int(a) if a else None for a in arenas])
        _check(lib().stage_gpu_exchange_attach_ptrs(self.h, C.cast(arr, _vp)))

    def arena(self):
        return int(lib().stage_gpu_exchange_arena(self.h))

    def post(self, splitters, keys_dev, n, flags=0, cuda_stream=0):
        sp = None
        if self.world > 1:
            sp = np.ascontiguousarray(splitters, np.uint8).reshape(self.world - 1, self.schema.key_size)
        _check(lib().stage_gpu_exchange_post(self.h, _ptr(sp), _ptr(keys_dev), n, flags, _vp(int(cuda_stream))))

    def serve(self, shard, read_ts, flags=0, cuda_stream=0):
        _check(lib().stage_gpu_exchange_serve(self.h, shard.h, read_ts, flags, _vp(int(cuda_stream))))

    def serve_project(self, shard, read_ts, col_off, col_len, flags=0, cuda_stream=0):
        _check(lib().stage_gpu_exchange_serve_project(self.h, shard.h, read_ts, col_off, col_len, flags, _vp(int(cuda_stream))))

    def serve_aggregate_i32(self, shard, read_ts, col_off, threshold, flags=0, cuda_stream=0):
        _check(lib().stage_gpu_exchange_serve_aggregate_i32(self.h, shard.h, read_ts, col_off, threshold, flags,
                                                            _vp(int(cuda_stream))))

    def aggregate_result(self, cuda_stream=0):
        out = AggI32()
        _check(lib().stage_gpu_exchange_aggregate_result(self.h, C.byref(out), _vp(int(cuda_stream))))
        return out

    def project_results_torch(self, n, col_len):
        """(found u8[n], ts i32[n], col u8[n, col_len]) torch views of the result slab after serve_project."""
        import torch
        f, t, c, st = _vp(), _vp(), _vp(), _u32(0)
        _check(lib().stage_gpu_exchange_project_results(self.h, C.byref(f), C.byref(t), C.byref(c), C.byref(st)))
        dev = torch.device("cuda", torch.cuda.current_device())
        found = torch.as_tensor(_DeviceView(int(f.value), (n,), "|u1", self), device=dev)
        ts = torch.as_tensor(_DeviceView(int(t.value), (n,), "<i4", self), device=dev)
        col = torch.as_tensor(_DeviceView(int(c.value), (n, int(st.value)), "|u1", self), device=dev)[:, :col_len]
        return found, ts, col

    def post_scans(self, splitters, start_keys_dev, scan_sizes_dev, n, flags=0, cuda_stream=0):
        sp = None
        if self.world > 1:
            sp = np.ascontiguousarray(splitters, np.uint8).reshape(self.world - 1, self.schema.key_size)
        _check(lib().stage_gpu_exchange_post_scans(self.h, _ptr(sp), _ptr(start_keys_dev), _ptr(scan_sizes_dev), n, flags,
                                                   _vp(int(cuda_stream))))

    def serve_scans(self, shard, read_ts, apply_visibility=True, flags=0, cuda_stream=0):
        _check(lib().stage_gpu_exchange_serve_scans(self.h, shard.h, read_ts, int(apply_visibility), flags,
                                                    _vp(int(cuda_stream))))

    def wait(self, cuda_stream=0):
        _check(lib().stage_gpu_exchange_wait(self.h, _vp(int(cuda_stream))))

    def peer_copy(self, dst_rank, dst_offset, nbytes, cuda_stream=0):
        _check(lib().stage_gpu_exchange_peer_copy(self.h, dst_rank, dst_offset, nbytes, _vp(int(cuda_stream))))

    def error(self, cuda_stream=0):
        e = _u32(0)
        _check(lib().stage_gpu_exchange_error(self.h, _vp(int(cuda_stream)), C.byref(e)))
        return int(e.value)

    def enable_timing(self, on=True):
        _check(lib().stage_gpu_exchange_enable_timing(self.h, int(on)))

    def last_timing(self):
        """(post_ms, search_ms, rows_ms, wait_ms) of the last post/serve/wait sequence."""
        ms = (C.c_float * 4)()
        _check(lib().stage_gpu_exchange_last_timing(self.h, ms))
        return tuple(float(v) for v in ms)

    def result_ptrs(self):
        f, t, p, st = _vp(), _vp(), _vp(), _u32(0)
        _check(lib().stage_gpu_exchange_results(self.h, C.byref(f), C.byref(t), C.byref(p), C.byref(st)))
        return int(f.value), int(t.value), int(p.value), int(st.value)

    def results_torch(self, n=None):
        """(found u8[n], ts i32[n], payload u8[n,P]) torch views of the result slab (no copy; payload rows are
        strided by the slab's padded row stride)."""
        import torch
        n = self.capacity if n is None else n
        f, t, p, stride = self.result_ptrs()
        P = self.schema.payload_size
        dev = torch.device("cuda", torch.cuda.current_device())
        found = torch.as_tensor(_DeviceView(f, (n,), "|u1", self), device=dev)
        ts = torch.as_tensor(_DeviceView(t, (n,), "<i4", self), device=dev)
        pay = torch.as_tensor(_DeviceView(p, (n, stride), "|u1", self), device=dev)[:, :P]
        return found, ts, pay

    def scan_results_torch(self, n=None):
        """(row_counts i32[n], keys u8[n,R,K], ts i32[n,R], payload u8[n,R,P]) torch views of the scan result slab
        (no copy; R = rows per scan of this exchange; rows beyond row_counts[q] are stale)."""
        import torch
        n = self.capacity if n is None else n
        c, k, t, p = _vp(), _vp(), _vp(), _vp()
        R, ks, ps = _u32(0), _u32(0), _u32(0)
        _check(lib().stage_gpu_exchange_scan_results(self.h, C.byref(c), C.byref(k), C.byref(t), C.byref(p), C.byref(R),
                                                     C.byref(ks), C.byref(ps)))
        R, ks, ps = int(R.value), int(ks.value), int(ps.value)
        K, P = self.schema.key_size, self.schema.payload_size
        dev = torch.device("cuda", torch.cuda.current_device())
        counts = torch.as_tensor(_DeviceView(int(c.value), (n,), "<i4", self), device=dev)
        keys = torch.as_tensor(_DeviceView(int(k.value), (n, R, ks), "|u1", self), device=dev)[:, :, :K]
        ts = torch.as_tensor(_DeviceView(int(t.value), (n, R), "<i4", self), device=dev)
        pay = torch.as_tensor(_DeviceView(int(p.value), (n, R, ps), "|u1", self), device=dev)[:, :, :P]
        return counts, keys, ts, pay

    def close(self):
        if self.h:
            _check(lib().stage_gpu_exchange_destroy(self.h))
            self.h = None
