"""GPU: stage_gpu_probe_aggregate_i32_noalloc — the allocation-free variant of the aggregate push-down (scratch in a fixed
device array inside the library, include/stage_gpu.h). Called through the C ABI (ctypes) with device buffers; its record
must equal the one stage_gpu_probe_aggregate_i32 returns for the same batch — integer work, so bit-exact like every other
test here — on two calls in a row (the second call finds the first call's leftovers in the shared scratch: a larger
batch first, then a smaller one with other keys), plus the size limit and an empty batch.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

import stage_indexorganized_b200 as sg

pytestmark = pytest.mark.gpu

MAX_PROBES = 1 << 20
P = 200


def _record(dev_out):
    import torch
    torch.cuda.synchronize()
    raw = dev_out.cpu().numpy().tobytes()
    found, below = np.frombuffer(raw, "<u8", 2, 0)
    mn, mx = np.frombuffer(raw, "<i4", 2, 16)
    return int(found), int(below), int(mn), int(mx)


def _noalloc(snap, d_keys, n, read_ts, col_off, thr, d_out):
    rc = sg.lib().stage_gpu_probe_aggregate_i32_noalloc(snap.h, C.c_void_p(d_keys.data_ptr()), n, read_ts, col_off, thr,
                                                        C.c_void_p(d_out.data_ptr()))
    return rc


def test_noalloc_aggregate_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    sg.init(0)
    n_rows = 300_000
    with sg.Snapshot.generate_ycsb(n_rows, P) as snap:
        snap.set_stream(torch.cuda.current_stream().cuda_stream)
        # give the column something to reduce: committed updates put other bytes (and version chains) into some rows
        rng = np.random.default_rng(123)
        upd = np.unique(rng.integers(0, n_rows, 20_000).astype(np.uint32))
        patches = rng.integers(0, 256, size=(len(upd), 8), dtype=np.uint8)
        assert snap.apply_updates(upd.view(np.uint8).reshape(-1, 4), patches, np.arange(100, 100 + len(upd), dtype=np.uint32)) == len(upd)
        d_out = torch.zeros(24, dtype=torch.uint8, device="cuda")
        batches = [(600_000, 11), (50_000, 12)]           # larger first: the second call sees stale scratch beyond its n
        for n, seed in batches:
            r = np.random.default_rng(seed)
            keys = r.integers(0, n_rows + n_rows // 10, n).astype(np.uint32)      # ~9 % misses
            d_keys = torch.from_numpy(keys.view(np.int32)).cuda()
            for read_ts, col_off, thr in ((50, 0, 0), (1_000_000, 0, 1 << 30), (5000, 3, -5)):
                want = snap.probe_aggregate_i32(d_keys, read_ts, col_off, thr, n=n)
                assert _noalloc(snap, d_keys, n, read_ts, col_off, thr, d_out) == 0
                got = _record(d_out)
                assert got == (want.found, want.below, want.min_value, want.max_value), (n, read_ts, col_off, thr)
                assert got[0] > 0
        # same batch twice, back to back without a synchronisation in between (same stream: ordered)
        keys = np.random.default_rng(13).integers(0, n_rows, 200_000).astype(np.uint32)
        d_keys = torch.from_numpy(keys.view(np.int32)).cuda()
        want = snap.probe_aggregate_i32(d_keys, 777, 0, 0, n=len(keys))
        d_out2 = torch.zeros(24, dtype=torch.uint8, device="cuda")
        assert _noalloc(snap, d_keys, len(keys), 777, 0, 0, d_out) == 0
        assert _noalloc(snap, d_keys, len(keys), 777, 0, 0, d_out2) == 0
        assert _record(d_out) == _record(d_out2) == (want.found, want.below, want.min_value, want.max_value)


def test_noalloc_limits_and_empty_batch():
    import torch
    sg.init(0)
    with sg.Snapshot.generate_ycsb(10_000, P) as snap:
        snap.set_stream(torch.cuda.current_stream().cuda_stream)
        d_out = torch.full((24,), 0xAB, dtype=torch.uint8, device="cuda")
        d_keys = torch.zeros(16, dtype=torch.int32, device="cuda")
        # beyond the fixed scratch: refused with a message, nothing launched that could touch the record
        rc = _noalloc(snap, d_keys, MAX_PROBES + 1, 1, 0, 0, d_out)
        assert rc == -1 and b"exceed the fixed scratch" in sg.lib().stage_gpu_last_error()
        torch.cuda.synchronize()
        assert (d_out.cpu().numpy() == 0xAB).all()
        assert _noalloc(snap, d_keys, 16, 1, P - 3, 0, d_out) == -1               # column runs past the row
        # an empty batch gives the neutral record
        assert _noalloc(snap, d_keys, 0, 1, 0, 0, d_out) == 0
        assert _record(d_out) == (0, 0, 2**31 - 1, -2**31)
        # exactly the maximum is accepted
        keys = np.random.default_rng(3).integers(0, 10_050, MAX_PROBES).astype(np.uint32)
        d_big = torch.from_numpy(keys.view(np.int32)).cuda()
        want = snap.probe_aggregate_i32(d_big, 1, 0, 100, n=MAX_PROBES)
        assert _noalloc(snap, d_big, MAX_PROBES, 1, 0, 100, d_out) == 0
        assert _record(d_out) == (want.found, want.below, want.min_value, want.max_value)
