"""GPU: BASELINE configs[3] — TPC-C stock-level / CH-Q2* read pattern over the index-organized stock and order_line
tables at w = 30 (STAGE_TEST_TPCC_W overrides), bit-exact against the CPU checker — the unmodified reference when its
prebuilt library is there, else the C restatement — on the same host-built trees (rows from stage_indexorganized_b200/tpcc.py,
whose shapes tests/test_tpcc_rows.py pins to the reference's own BuildStockTuple / BuildOrderLineTuple).

Workload A (RunStockLevel, benchmark/tpcc/tpcc_stock_level.cpp:37-196): per transaction 20 MVCC scans of 10
order_line rows from {w, d, o, 5}, then point probes of stock for the OL_I_IDs of the rows of that order.
Workload B (CH-Q2*, benchmark/tpcc/tpcc_new_order.cpp:608-983): stock probes of one supplier's (w, i) pairs and
MIN(S_QUANTITY) over them. Both tables use TPC-C's ParameterSet (64 KiB split threshold, tpcc.cpp:78); their keys
are >= 16 bytes, i.e. memcmp-ordered raw little-endian integers."""
import os

import numpy as np
import pytest

import stage_indexorganized_b200 as sg
from stage_indexorganized_b200 import tpcc
from helpers import assert_probe_equal, assert_scan_equal
from oracle import pyoracle

pytestmark = pytest.mark.gpu
W = int(os.environ.get("STAGE_TEST_TPCC_W", "30"))


@pytest.fixture(scope="module")
def tables():
    sg.init(0)
    sk, sp = tpcc.stock_rows(W)
    ok, op, _ = tpcc.order_line_rows(W)
    if pyoracle.have_ref():        # the unmodified reference builds the trees (TPC-C's ParameterSet, tpcc.cpp:78) and checks the reads
        stock = pyoracle.RefTable(tpcc.STOCK_PAYLOAD, split_threshold=65536, catalog_key_len=16, leaf_pool_blocks=max(len(sk) // 40, 4096))
        ol = pyoracle.RefTable(tpcc.OL_PAYLOAD, split_threshold=65536, catalog_key_len=32, leaf_pool_blocks=max(len(ok) // 100, 4096))
    else:
        stock = pyoracle.OracleTable(tpcc.STOCK_PAYLOAD, split_threshold=65536)
        ol = pyoracle.OracleTable(tpcc.OL_PAYLOAD, split_threshold=65536)
    assert stock.insert(sk, sp, 0) == len(sk)
    assert ol.insert(ok, op, 0) == len(ok)
    # some committed updates of S_QUANTITY-bearing rows so that the MVCC branch is exercised too
    rng = np.random.default_rng(4)
    upd = sk[rng.integers(0, len(sk), 20000)]
    assert stock.commit_update_batch(upd, 10) == 20000
    s_snap = sg.Snapshot.from_tree(stock.root, tpcc.STOCK_KEY, tpcc.STOCK_PAYLOAD)
    o_snap = sg.Snapshot.from_tree(ol.root, tpcc.OL_KEY, tpcc.OL_PAYLOAD)
    yield stock, ol, s_snap, o_snap, len(sk), len(ok)
    s_snap.close()
    o_snap.close()


def test_table_sizes(tables):
    _, _, s_snap, o_snap, n_stock, n_ol = tables
    assert s_snap.info().n_rows == n_stock == W * tpcc.ITEMS
    assert o_snap.info().n_rows == n_ol
    assert s_snap.info().key_words == 4 and o_snap.info().key_words == 8


def test_stock_level_shape(tables):
    stock, ol, s_snap, o_snap, _, _ = tables
    start, w, d, o = tpcc.stock_level_requests(W, 2000)
    for read_ts in (1, 5000):
        got = o_snap.scan(start, 10, read_ts, 1)
        want = ol.scan(start, 10, read_ts, 1)
        assert_scan_equal(got, want, f"order_line scans ts={read_ts}")
    counts, keys, _, pay = got
    # the predicate of RunStockLevel keeps rows of the order asked for; their OL_I_IDs are probed in stock
    k64 = keys.reshape(len(start), 10, 32).view(np.int64).reshape(len(start), 10, 4)
    valid = np.arange(10)[None, :] < counts[:, None]
    match = valid & (k64[..., 0] == w[:, None]) & (k64[..., 1] == d[:, None]) & (k64[..., 2] == o[:, None])
    item = pay[..., 0:4].copy().view(np.int32).reshape(len(start), 10)
    ww = np.broadcast_to(w[:, None], match.shape)[match]
    probes = tpcc.stock_key(ww, item[match].astype(np.int64))
    assert len(probes) > 1000
    for read_ts in (1, 9, 10000, 30000):
        assert_probe_equal(s_snap.probe(probes, read_ts), stock.probe(probes, read_ts), f"stock probes ts={read_ts}")
    # the transaction's result: distinct items below the threshold — same rows, same answer
    f, _, _, p = s_snap.probe(probes, 30000)
    wf, _, _, wp = stock.probe(probes, 30000)
    q, wq = p[:, 0:4].copy().view(np.int32).ravel(), wp[:, 0:4].copy().view(np.int32).ravel()
    assert np.array_equal(np.unique(probes[(f == 1) & (q < 15)], axis=0), np.unique(probes[(wf == 1) & (wq < 15)], axis=0))


def test_q2_shape(tables):
    stock, _, s_snap, _, _, _ = tables
    for supplier in (0, 77, 4242, 9999):
        keys = tpcc.q2_supplier_keys(W, supplier)
        assert len(keys) > 0
        got, want = s_snap.probe(keys, 50000), stock.probe(keys, 50000)
        assert_probe_equal(got, want, f"supplier {supplier}")
        assert got[0].all()
        q = got[3][:, 0:4].copy().view(np.int32).ravel()
        wq = want[3][:, 0:4].copy().view(np.int32).ravel()
        assert q.min() == wq.min()


def test_stock_range_scans(tables):
    stock, _, s_snap, _, n_stock, _ = tables
    rng = np.random.default_rng(8)
    start = tpcc.stock_key(rng.integers(1, W + 1, 300), rng.integers(0, tpcc.ITEMS + 2, 300))
    for S in (10, 200):
        for vis, ts in ((0, 0), (1, 1), (1, 15000)):
            assert_scan_equal(s_snap.scan(start, S, ts, vis), stock.scan(start, S, ts, vis), f"S={S} vis={vis} ts={ts}")


def test_quantity_pushdown(tables):
    """Column push-down (stage_gpu_probe_project / scan_project): S_QUANTITY only — same rows, same bytes."""
    stock, ol, s_snap, o_snap, _, _ = tables
    keys = np.concatenate([tpcc.q2_supplier_keys(W, 1234), tpcc.stock_key([W + 5, 0], [1, 1])])     # + two misses
    for read_ts in (1, 50000):
        f, ts, col = s_snap.probe_column(keys, read_ts, 0, 4)
        wf, wts, _, wp = stock.probe(keys, read_ts)
        assert np.array_equal(f, wf) and np.array_equal(ts, wts)
        assert np.array_equal(col[wf == 1], wp[wf == 1][:, 0:4]) and not col[wf == 0].any()
        q = col[f == 1].copy().view(np.int32).ravel()
        assert q.min() == wp[wf == 1][:, 0:4].copy().view(np.int32).min()           # CH-Q2*: MIN(S_QUANTITY)
    # stock-level: OL_I_ID (payload bytes 0..3) of the 10 order_line rows per scan
    start, _, _, _ = tpcc.stock_level_requests(W, 500, seed=11)
    counts, kout, col = o_snap.scan_column(start, 10, 1, 1, 0, 4)
    wc, wk, _, wp = ol.scan(start, 10, 1, 1)
    assert np.array_equal(counts, wc) and np.array_equal(kout, wk)
    valid = np.arange(10)[None, :] < wc[:, None]
    assert np.array_equal(col[valid], wp[valid][:, 0:4]) and not col[~valid].any()


def test_quantity_aggregate_pushdown(tables):
    """Predicate / aggregate push-down (stage_gpu_probe_aggregate_i32): the stock-level count of S_QUANTITY < threshold
    and CH-Q2*'s MIN(S_QUANTITY), reduced on the device, against the checker's rows; host and device key buffers,
    an unaligned column and an empty batch."""
    import torch
    stock, _, s_snap, _, _, _ = tables
    rng = np.random.default_rng(21)
    keys = np.concatenate([tpcc.q2_supplier_keys(W, 77), tpcc.stock_key(rng.integers(1, W + 1, 5000), rng.integers(1, tpcc.ITEMS + 1, 5000)),
                           tpcc.stock_key([W + 3, 0, W + 9], [5, 1, 2])])                     # + three misses
    for read_ts in (1, 50000):
        wf, _, _, wp = stock.probe(keys, read_ts)
        wq = wp[wf == 1][:, 0:4].copy().view(np.int32).ravel()
        for thr in (15, 50, -1):
            a = s_snap.probe_aggregate_i32(keys, read_ts, 0, thr)
            assert (a.found, a.below, a.min_value, a.max_value) == (int(wf.sum()), int((wq < thr).sum()), int(wq.min()), int(wq.max()))
        dk = torch.from_numpy(keys).cuda()
        a = s_snap.probe_aggregate_i32(dk, read_ts, 0, 15, n=len(keys))
        assert (a.found, a.below, a.min_value) == (int(wf.sum()), int((wq < 15).sum()), int(wq.min()))
        # a column at an odd offset: bytes 3..6 read as a little-endian int32
        wo = np.ascontiguousarray(wp[wf == 1][:, 3:7]).view("<i4").ravel()
        a = s_snap.probe_aggregate_i32(keys, read_ts, 3, 0)
        assert (a.below, a.min_value, a.max_value) == (int((wo < 0).sum()), int(wo.min()), int(wo.max()))
    a = s_snap.probe_aggregate_i32(keys[:0], 1, 0, 15)
    assert (a.found, a.below, a.min_value, a.max_value) == (0, 0, 2**31 - 1, -2**31)
    a = s_snap.probe_aggregate_i32(tpcc.stock_key([W + 3], [5]), 1, 0, 15)                    # only a miss
    assert (a.found, a.min_value) == (0, 2**31 - 1)
    with pytest.raises(sg.StageError):
        s_snap.probe_aggregate_i32(keys, 1, tpcc.STOCK_PAYLOAD - 3, 15)                         # column runs past the row


def test_q2_group_aggregate_per_supplier(tables):
    """stage_gpu_probe_group_aggregate_i32: CH-Q2*'s loop over suppliers (tpcc_new_order.cpp:803-866) as ONE call — the
    stock probes of a run of suppliers, one record per supplier — against the reference's loop restated on the checker's
    rows: MIN(S_QUANTITY) with its first row, and the row the loop literally keeps (its running minimum is never
    lowered, so the last row below INT16_MAX wins). Groups with misses, empty groups, host and device buffers."""
    import torch
    stock, _, s_snap, _, _, _ = tables
    keys_all, off_all = tpcc.q2_all_supplier_keys(W)
    assert len(keys_all) == W * tpcc.ITEMS and off_all[-1] == len(keys_all)
    assert np.array_equal(keys_all[off_all[77]:off_all[78]], tpcc.q2_supplier_keys(W, 77))
    g0, g1 = 4200, 4264                                       # 64 suppliers
    lo, hi = int(off_all[g0]), int(off_all[g1])
    keys = np.concatenate([keys_all[lo:hi], tpcc.stock_key([W + 3, 0, W + 9], [5, 1, 2])])      # + a group of misses
    off = np.concatenate([off_all[g0:g1 + 1] - lo, [hi - lo, hi - lo + 3]]).astype(np.uint64)   # + an empty group
    for read_ts in (1, 50000):
        wf, _, _, wp = stock.probe(keys, read_ts)
        q = np.ascontiguousarray(wp[:, 0:4]).view("<i4").ravel()
        for thr in (32767, 40):
            got = s_snap.probe_group_aggregate_i32(keys, off, read_ts, 0, thr)
            assert len(got) == len(off) - 1
            for g in range(len(off) - 1):
                a, b = int(off[g]), int(off[g + 1])
                hit = np.flatnonzero(wf[a:b] == 1) + a
                r = got[g]
                if len(hit) == 0:
                    assert (r["found"], r["below"], r["min_value"], r["min_index"], r["max_value"], r["last_below_index"]) == \
                        (0, 0, 2**31 - 1, 0xFFFFFFFF, -2**31, 0xFFFFFFFF), g
                    continue
                qq = q[hit]
                below = hit[qq < thr]
                keep = 0xFFFFFFFF                              # the reference's loop, literally
                for j in hit:
                    if thr > q[j]:
                        keep = j
                assert (r["found"], r["below"], r["min_value"], r["max_value"]) == (len(hit), len(below), qq.min(), qq.max()), g
                assert r["min_index"] == hit[np.argmin(qq)] and r["last_below_index"] == keep, g
        dk, do = torch.from_numpy(keys).cuda(), torch.from_numpy(off.astype(np.int64)).cuda()
        dout = torch.zeros((len(off) - 1, 24), dtype=torch.uint8, device="cuda")
        s_snap.probe_group_aggregate_i32(dk, do, read_ts, 0, 40, n=len(keys), out=dout)
        torch.cuda.synchronize()
        assert np.array_equal(dout.cpu().numpy().view(sg.GROUP_AGG_I32).ravel(), s_snap.probe_group_aggregate_i32(keys, off, read_ts, 0, 40))
    with pytest.raises(sg.StageError):
        s_snap.probe_group_aggregate_i32(keys, off[:-1], 1, 0, 40)                             # offsets do not cover the batch
    with pytest.raises(sg.StageError):
        s_snap.probe_group_aggregate_i32(keys, off[::-1].copy(), 1, 0, 40)                     # not ascending


def _stock_level_by_the_three_steps(ol, stock, reqs, read_ts):
    """RunStockLevel's reads done the reference's way on the checker's tables (tpcc_stock_level.cpp:92-190): 20 scans of
    10 rows, first returned row of the order -> one stock probe -> S_QUANTITY < threshold -> distinct items."""
    n = len(reqs)
    w = np.repeat(reqs["w_id"], 20)
    d = np.repeat(reqs["d_id"], 20)
    o = (reqs["next_o_id"][:, None] - 20 + np.arange(20)[None, :]).ravel()
    start = tpcc.order_line_key(w, d, o, np.full(len(w), 5, np.int64))
    counts, keys, _, pay = ol.scan(start, 10, read_ts, 1)
    k64 = keys.reshape(len(start), 10, 32).view(np.int64).reshape(len(start), 10, 4)
    valid = np.arange(10)[None, :] < counts[:, None]
    match = valid & (k64[..., 0] == w[:, None]) & (k64[..., 1] == d[:, None]) & (k64[..., 2] == o[:, None])
    has = match.any(axis=1)
    first = match.argmax(axis=1)
    item = pay[..., 0:4].copy().view(np.int32).reshape(len(start), 10)[np.arange(len(start)), first].astype(np.int64)
    probes = tpcc.stock_key(w[has], item[has])
    f, _, _, p = stock.probe(probes, read_ts)
    qty = p[:, 0:4].copy().view(np.int32).ravel()
    thr = np.repeat(reqs["threshold"], 20)[has]
    below = (f == 1) & (qty < thr)
    txn = np.repeat(np.arange(n), 20)[has]
    out = np.zeros(n, np.uint32)
    for t in np.unique(txn[below]):
        out[t] = len(np.unique(item[has][below & (txn == t)]))
    return out


def test_stock_level_operator(tables):
    """stage_gpu_tpcc_stock_level: the whole read side of RunStockLevel in one call, counts only coming back — against the
    three steps run through the checker. Requests include districts' last orders, order ids below the first order
    (scans that find nothing of the order) and thresholds from "nothing" to "everything"."""
    import torch
    stock, ol, s_snap, o_snap, _, _ = tables
    rng = np.random.default_rng(31)
    n = 3000
    reqs = np.zeros(n, sg.STOCK_LEVEL_REQ)
    reqs["w_id"] = rng.integers(1, W + 1, n)
    reqs["d_id"] = rng.integers(1, tpcc.DISTRICTS + 1, n)
    reqs["next_o_id"] = rng.integers(21, tpcc.ORDERS + 2, n)
    reqs["threshold"] = rng.integers(10, 21, n)
    reqs["next_o_id"][:50] = tpcc.ORDERS + 1                    # the district's newest orders (what the driver asks for)
    reqs["next_o_id"][50:60] = 5                                # order ids -15 .. 4: nothing (or little) to find
    reqs["threshold"][60:80] = 1000                             # every probed item counts
    reqs["threshold"][80:100] = 0                               # none does
    for read_ts in (1, 9000, 40000):
        want = _stock_level_by_the_three_steps(ol, stock, reqs, read_ts)
        got = sg.tpcc_stock_level(o_snap, s_snap, reqs, read_ts)
        assert np.array_equal(got, want), (read_ts, np.nonzero(got != want)[0][:10], got[:10], want[:10])
        assert want[60:80].max() > 10 and not want[80:100].any()
    # device-resident requests / results
    d_req = torch.from_numpy(reqs.view(np.uint8).reshape(n, 32)).cuda()
    d_out = torch.zeros(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    sg.tpcc_stock_level(o_snap, s_snap, d_req, 40000, out=d_out, n=n)
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy().view(np.uint32), _stock_level_by_the_three_steps(ol, stock, reqs, 40000))
    with pytest.raises(sg.StageError):
        sg.tpcc_stock_level(s_snap, o_snap, reqs, 1)             # tables swapped: key sizes do not fit
