"""Closed-form model of the device-generated YCSB int table (csrc/generate.cuh) and of the committed updates installed
with stage_gpu_snapshot_apply_updates — what a reader must see, computed without the table.

Row function (benchmark/ycsb/ycsb_loader.cpp:144-159 of the reference): key = uint32 rowid, payload = payload_size x
(uint8) rowid, commit id 0. Update i of a batch rewrites payload bytes [100, 200) with (uint8) commit id and commits
with id FIRST_CID + i (the shape of the YCSB update transaction after commit, ycsb_mixed.cpp:38-80). Visibility follows
the executors: point probes (include/execute/executor.h:383-446) take the in-leaf version if read_ts >= its commit id,
else the newest retired version with begin <= read_ts <= commit, where the loader's version (begin 0) ends the walk;
range scans (executor.h:465-534) take the in-leaf version or drop the row.

Used by tests/test_gpu_fullsize.py (where it is pinned against the CPU checker at a size the checker handles) and by
bench.py's parity_check at N > 1, where no host tree of the 125 M-row shards can exist.
"""
import numpy as np

FIRST_CID = 10


class UpdateModel:
    """Committed updates (key, commit id = first_cid + i) -> what a reader sees."""

    def __init__(self, upd_keys, first_cid=FIRST_CID):
        k = np.asarray(upd_keys, np.uint64)
        cid = np.arange(first_cid, first_cid + len(k), dtype=np.uint64)
        self.code = np.sort((k << np.uint64(32)) | cid)           # (key, cid) pairs, sorted
        self.keys_sorted = self.code >> np.uint64(32)

    def visible(self, keys, read_ts, scan=False):
        """-> (visible bool[n], stamp u32[n]) for existing keys. Point probes walk the version chain (an older
        version with begin != 0 can answer); a range scan only ever returns the in-leaf version, i.e. the row is
        visible iff ALL its updates committed at or before read_ts (executor.h:465-534 with next_tuple_ptr == 0)."""
        k = np.asarray(keys, np.uint64)
        lo = np.searchsorted(self.keys_sorted, k, "left")
        hi = np.searchsorted(self.keys_sorted, k, "right")
        n_upd = hi - lo
        # number of this key's updates with cid <= read_ts
        le = np.searchsorted(self.code, (k << np.uint64(32)) | np.uint64(min(read_ts, 0xFFFFFFFF)), "right") - lo
        le = np.clip(le, 0, n_upd)
        stamp = np.where(le > 0, self.code[np.clip(lo + le - 1, 0, max(len(self.code) - 1, 0))] & np.uint64(0xFFFFFFFF), 0)
        vis = (n_upd == 0) | ((le == n_upd) if scan else (le > 0))
        return vis, np.where(vis, stamp, 0).astype(np.uint32)


def expected_payload_bytes(keys, stamp):
    """(first byte, byte 150, last byte) of the row image: base = (uint8) key, column [100,200) = (uint8) stamp if updated."""
    base = (np.asarray(keys, np.uint32) & 0xFF).astype(np.uint8)
    col = np.where(np.asarray(stamp) > 0, (np.asarray(stamp) & 0xFF).astype(np.uint8), base)
    return base, col, base
