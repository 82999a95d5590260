"""Seeded restatement of the TPC-C rows the read path touches (BASELINE configs[3]: stock-level / CH-Q2*).

Key and tuple shapes follow the reference's benchmark/tpcc/tpcc_record.h:
  Stock      key {int64 S_W_ID, int64 S_I_ID} (16 B, :417-430), tuple 416 B => payload 400 B
             [S_QUANTITY i32][S_YTD i32][S_ORDER_CNT i32][S_REMOTE_CNT i32][S_DIST_01..10 32 B each][S_DATA 64 B]
  OrderLine  key {int64 OL_W_ID, OL_D_ID, OL_O_ID, OL_NUMBER} (32 B, :694-714), payload 60 B
             [OL_I_ID i32][OL_SUPPLY_W_ID i32][OL_DELIVERY_D i64][OL_QUANTITY i32][OL_AMOUNT 8 B][OL_DIST_INFO 32 B]
Field values follow BuildStockTuple / BuildOrderLineTuple (benchmark/tpcc/tpcc_loader.cpp:1281-1374) with a
seeded numpy generator in place of the loader's un-seeded rand() (SURVEY.md §8d, config 4 input): the loader's
"random strings" are ONE random alphanumeric character repeated (GetRandomAlphaNumericString, tpcc_loader.cpp:849-860),
S_DIST_xx = 31 of them + NUL (name_length = 31, strncpy into char[32]), S_DATA = 64, OL_DIST_INFO = 32.
tests/test_tpcc_rows.py holds these rows to the bytes the reference's own builders produce (oracle/_ref).
Both keys are >= 16 bytes, so the reference orders them with memcmp over the raw little-endian bytes
(include/vstore/b_tree.h:116-134) — NOT numerically; the table order is whatever that gives.
"""
import numpy as np

STOCK_KEY, STOCK_PAYLOAD = 16, 400
OL_KEY, OL_PAYLOAD = 32, 60
ITEMS, DISTRICTS, ORDERS = 100_000, 10, 3000
# GetRandomAlphaNumericString draws an index in [0, sizeof(alphanumeric) - 1], i.e. the terminating NUL included
# (tpcc_loader.cpp:850-857): 1 string in 63 is all zero bytes
_ALNUM = np.frombuffer(b"0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz\x00", np.uint8)


def stock_key(w, i):
    k = np.empty((len(np.atleast_1d(w)), 2), np.int64)
    k[:, 0], k[:, 1] = w, i
    return k.view(np.uint8).reshape(-1, STOCK_KEY)


def order_line_key(w, d, o, n):
    k = np.empty((len(np.atleast_1d(w)), 4), np.int64)
    k[:, 0], k[:, 1], k[:, 2], k[:, 3] = w, d, o, n
    return k.view(np.uint8).reshape(-1, OL_KEY)


def stock_rows(warehouses, items=ITEMS, seed=1):
    """(keys [n,16], payloads [n,400]) in loader order: warehouse-major, item-minor (tpcc_loader.cpp LoadWarehouses)."""
    rng = np.random.default_rng(seed)
    w = np.repeat(np.arange(1, warehouses + 1, dtype=np.int64), items)
    i = np.tile(np.arange(1, items + 1, dtype=np.int64), warehouses)
    n = len(w)
    pay = np.zeros((n, STOCK_PAYLOAD), np.uint8)
    pay[:, 0:4] = rng.integers(10, 101, n).astype(np.int32).view(np.uint8).reshape(n, 4)      # S_QUANTITY 10..100
    dist = np.repeat(_ALNUM[rng.integers(0, len(_ALNUM), size=n)][:, None], 32, axis=1)         # one character, 31 times
    dist[:, 31] = 0
    for c in range(10):                                                                        # the same string ten times
        pay[:, 16 + 32 * c: 48 + 32 * c] = dist
    pay[:, 336:400] = _ALNUM[rng.integers(0, len(_ALNUM), size=n)][:, None]                     # S_DATA: one character x 64
    return stock_key(w, i), pay


def order_line_rows(warehouses, districts=DISTRICTS, orders=ORDERS, seed=2):
    """(keys [n,32], payloads [n,60], OL_I_ID [n]) in loader order (w, d, o, ol_number), 5..15 lines per order."""
    rng = np.random.default_rng(seed)
    n_orders = warehouses * districts * orders
    cnt = rng.integers(5, 16, n_orders)
    total = int(cnt.sum())
    oid = np.repeat(np.arange(n_orders, dtype=np.int64), cnt)
    first = np.cumsum(cnt) - cnt
    number = np.arange(total, dtype=np.int64) - np.repeat(first, cnt) + 1
    w = oid // (districts * orders) + 1
    d = (oid // orders) % districts + 1
    o = oid % orders + 1
    item = rng.integers(0, ITEMS + 1, total).astype(np.int32)
    pay = np.zeros((total, OL_PAYLOAD), np.uint8)
    pay[:, 0:4] = item.view(np.uint8).reshape(total, 4)
    pay[:, 4:8] = w.astype(np.int32).view(np.uint8).reshape(total, 4)
    pay[:, 8:16] = (1_600_000_000 + oid).view(np.uint8).reshape(total, 8)
    pay[:, 16:20] = np.full(total, 5, np.int32).view(np.uint8).reshape(total, 4)
    pay[:, 20:28] = np.frombuffer(b"0.000000", np.uint8)
    pay[:, 28:60] = _ALNUM[rng.integers(0, len(_ALNUM), size=total)][:, None]                   # OL_DIST_INFO: one character x 32
    return order_line_key(w, d, o, number), pay, item


def stock_level_requests(warehouses, txns, seed=3, districts=DISTRICTS, orders=ORDERS):
    """RunStockLevel's read pattern (benchmark/tpcc/tpcc_stock_level.cpp:37-196): per transaction a random (w, d) and,
    for the 20 orders below the district's next order id, one order_line scan of 10 rows starting at
    {w, d, o, 5}. Returns (start_keys [txns*20, 32], w [txns*20], d, o)."""
    rng = np.random.default_rng(seed)
    w = np.repeat(rng.integers(1, warehouses + 1, txns), 20).astype(np.int64)
    d = np.repeat(rng.integers(1, districts + 1, txns), 20).astype(np.int64)
    o = np.tile(np.arange(orders + 1 - 20, orders + 1, dtype=np.int64), txns)
    return order_line_key(w, d, o, np.full(len(w), 5, np.int64)), w, d, o


def q2_supplier_keys(warehouses, supplier, items=ITEMS):
    """CH-Q2*'s stock probes for one supplier: every (w, i) with (w * i) mod 10000 == supplier
    (benchmark/tpcc/tpcc_workload.cpp:581-585) — about warehouses * items / 10000 keys."""
    w = np.repeat(np.arange(1, warehouses + 1, dtype=np.int64), items)
    i = np.tile(np.arange(1, items + 1, dtype=np.int64), warehouses)
    m = (w * i) % 10000 == supplier
    return stock_key(w[m], i[m])


def q2_all_supplier_keys(warehouses, items=ITEMS, suppliers=10000):
    """supp_stock_map (benchmark/tpcc/tpcc_workload.cpp:581-585) as one probe batch: the (w, i) pairs grouped by
    (w * i) mod 10000, within a supplier in the map's order (w ascending, then i ascending).
    Returns (keys [warehouses * items, 16], group_offsets uint64 [suppliers + 1])."""
    w = np.repeat(np.arange(1, warehouses + 1, dtype=np.int64), items)
    i = np.tile(np.arange(1, items + 1, dtype=np.int64), warehouses)
    sup = (w * i) % suppliers
    order = np.argsort(sup, kind="stable")
    off = np.zeros(suppliers + 1, np.uint64)
    off[1:] = np.cumsum(np.bincount(sup, minlength=suppliers))
    return stock_key(w[order], i[order]), off
