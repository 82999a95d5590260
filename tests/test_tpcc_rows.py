"""The package's seeded TPC-C row functions (stage_indexorganized_b200/tpcc.py) against the reference's own builders
(BuildStockTuple / BuildOrderLineTuple, /root/reference/benchmark/tpcc/tpcc_loader.cpp:1281-1374, compiled unmodified
into oracle/_ref/libstage_ref_tpcc.so): identical key bytes and sizes, identical deterministic fields, and every
randomised field inside the shape the reference produces (checked by one validator applied to both sides)."""
import ctypes as C
import os

import numpy as np
import pytest

from stage_indexorganized_b200 import tpcc

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libstage_ref_tpcc.so")
pytestmark = pytest.mark.skipif(not os.path.exists(LIB), reason="oracle/_ref/libstage_ref_tpcc.so not built")
_ALNUM = set(b"0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz\x00")    # the loader can draw the NUL too


@pytest.fixture(scope="module")
def ref():
    l = C.CDLL(LIB)
    l.ref_tpcc_config(30, tpcc.ITEMS)
    return l


def _one_char(b, n):
    return len(b) == n and b[0] in _ALNUM and all(x == b[0] for x in b)


def stock_shape_ok(pay):
    q, ytd, oc, rc = pay[:16].view(np.int32)
    d = [bytes(pay[16 + 32 * c: 48 + 32 * c]) for c in range(10)]
    return (10 <= q <= 100 and ytd == 0 and oc == 0 and rc == 0 and all(x == d[0] for x in d)
            and _one_char(d[0][:31], 31) and d[0][31] == 0 and _one_char(bytes(pay[336:400]), 64))


def order_line_shape_ok(pay, w):
    item, supply = pay[:8].view(np.int32)
    return (0 <= item <= tpcc.ITEMS and supply == w and pay[16:20].view(np.int32)[0] == 5
            and bytes(pay[20:28]) == b"0.000000" and _one_char(bytes(pay[28:60]), 32))


def test_sizes_match_reference_structs(ref):
    k, p = C.c_uint32(0), C.c_uint32(0)
    ref.ref_tpcc_stock_sizes(C.byref(k), C.byref(p))
    assert (k.value, p.value) == (tpcc.STOCK_KEY, tpcc.STOCK_PAYLOAD)
    ref.ref_tpcc_order_line_sizes(C.byref(k), C.byref(p))
    assert (k.value, p.value) == (tpcc.OL_KEY, tpcc.OL_PAYLOAD)


def test_stock_rows_have_the_reference_shape(ref):
    keys, pays = tpcc.stock_rows(2, items=500, seed=5)
    rng = np.random.default_rng(0)
    for idx in rng.integers(0, len(keys), 200):
        w, i = int(idx // 500 + 1), int(idx % 500 + 1)
        rk, rp = np.zeros(tpcc.STOCK_KEY, np.uint8), np.zeros(tpcc.STOCK_PAYLOAD, np.uint8)
        ref.ref_tpcc_stock_tuple(i, w, rk.ctypes.data_as(C.c_void_p), rp.ctypes.data_as(C.c_void_p))
        assert np.array_equal(keys[idx], rk)                       # the key bytes the reference inserts under
        assert stock_shape_ok(rp), "validator rejects a reference row"
        assert stock_shape_ok(pays[idx]), f"restated stock row {idx} is outside the reference's shape"


def test_order_line_rows_have_the_reference_shape(ref):
    keys, pays, items = tpcc.order_line_rows(1, districts=2, orders=40, seed=6)
    k64 = keys.view(np.int64).reshape(-1, 4)
    rng = np.random.default_rng(1)
    for idx in rng.integers(0, len(keys), 200):
        w, d, o, n = (int(x) for x in k64[idx])
        rk, rp = np.zeros(tpcc.OL_KEY, np.uint8), np.zeros(tpcc.OL_PAYLOAD, np.uint8)
        ref.ref_tpcc_order_line_tuple(o, d, w, n, w, 0, rk.ctypes.data_as(C.c_void_p), rp.ctypes.data_as(C.c_void_p))
        assert np.array_equal(keys[idx], rk)
        assert order_line_shape_ok(rp, w), "validator rejects a reference row"
        assert order_line_shape_ok(pays[idx], w), f"restated order_line row {idx} is outside the reference's shape"
        assert pays[idx][:4].view(np.int32)[0] == items[idx]
