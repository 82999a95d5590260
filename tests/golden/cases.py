"""Deterministic table/key-stream recipes shared by make_golden.py (which runs them on the live
reference) and the tests (which run them on the C oracle and on the CUDA path)."""
import zlib

import numpy as np

from oracle import pyoracle


def u32_keys(a):
    return np.ascontiguousarray(np.asarray(a, np.uint32)).view(np.uint8).reshape(-1, 4)


def row_crc(keys, pays, valid):
    """crc32 over key||payload of every valid row (0 elsewhere) — keeps the fixtures small."""
    flat_k = keys.reshape(-1, keys.shape[-1])
    flat_p = pays.reshape(-1, pays.shape[-1])
    v = np.asarray(valid).reshape(-1)
    out = np.zeros(flat_k.shape[0], np.uint32)
    for i in np.flatnonzero(v):
        out[i] = zlib.crc32(flat_p[i].tobytes(), zlib.crc32(flat_k[i].tobytes()))
    return out.reshape(np.asarray(valid).shape)


def scan_valid(counts, stride):
    return np.arange(stride)[None, :] < counts[:, None]


def case_ycsb10k(cls):
    """BASELINE config #1 table: 10 000-row YCSB int table, single loader thread."""
    t = cls(1000)
    assert t.load_ycsb_int(0, 10000) == 10000
    probes = np.concatenate([pyoracle.uniform_keys(10000, 4096, seed=1),
                             pyoracle.Zipf(9999, 0.2, seed=1).draw(2048).astype(np.uint32),
                             np.array([10000, 10001, 0xFFFFFFFF, 0x80808080, 70000], np.uint32)])
    scans = pyoracle.uniform_keys(10000, 64, seed=3)
    return dict(table=t, key_size=4, payload=1000, leaf=65536, probes=u32_keys(probes), probe_ts=[1],
                scans=u32_keys(scans), scan_sizes=[10, 1000], scan_ts=[1])


def case_mvcc(cls):
    """6000-row table, 1500 committed updates (chains, some rows updated several times), 300 deletes."""
    N, P = 6000, 1000
    t = cls(P)
    assert t.load_ycsb_int(0, N) == N
    rng = np.random.default_rng(11)
    upd = rng.integers(0, N, 1500).astype(np.uint32)
    assert t.commit_update_batch(u32_keys(upd), 10) == 1500
    for k in rng.choice(N, 300, replace=False):
        t.delete(u32_keys([k])[0])
    probes = rng.integers(0, N + 10, 3000).astype(np.uint32)
    scans = rng.integers(0, N, 48).astype(np.uint32)
    return dict(table=t, key_size=4, payload=P, leaf=65536, probes=u32_keys(probes),
                probe_ts=[0, 9, 10, 700, 1509, 1510, 4000], scans=u32_keys(scans), scan_sizes=[10, 200],
                scan_ts=[0, 700, 4000])


def case_string16(cls):
    """16-byte memcmp-ordered keys (YCSB string mode: FastRandom.next_string(16),
    benchmark/ycsb/ycsb_loader.cpp:107-117; TPC-C stock has the same key/compare shape), payload 400."""
    n, P = 4000, 400
    keys = pyoracle.random_strings(1804289383, n, 16)
    pay = np.empty((n, P), np.uint8)
    pay[:] = np.arange(n, dtype=np.uint32)[:, None].astype(np.uint8)
    pay[:, :16] = keys
    t = cls(P, split_threshold=65536, catalog_key_len=16) if cls is pyoracle.RefTable else cls(P, split_threshold=65536)
    assert t.insert(keys, pay, 0) == n
    rng = np.random.default_rng(5)
    probes = np.concatenate([keys[rng.integers(0, n, 2000)], rng.integers(0, 256, size=(200, 16), dtype=np.uint8)])
    return dict(table=t, key_size=16, payload=P, leaf=65536, probes=probes, probe_ts=[1], scans=probes[:48],
                scan_sizes=[10, 300], scan_ts=[1])


def case_btree_kat(cls):
    """test/testing_btree.cpp fixture: ParameterSet(3072,1024,4096,8), keys std::to_string(i) for
    i = 1000..9999 (all 4 bytes), payload = i as 8 bytes (RangeScanBySize KAT, :630-675)."""
    t = cls(8, split_threshold=3072, merge_threshold=1024, leaf_node_size=4096)
    ids = np.arange(1000, 10000, dtype=np.uint64)
    keys = np.array([list(str(i).encode()) for i in ids], np.uint8)
    assert t.insert(keys, ids.view(np.uint8).reshape(-1, 8), 0) == len(ids)
    scans = np.array([list(b"9000"), list(b"1000"), list(b"9999"), list(b"5555"), list(b"0000")], np.uint8)
    return dict(table=t, key_size=4, payload=8, leaf=4096, probes=keys[::7], probe_ts=[0], scans=scans,
                scan_sizes=[100, 1000], scan_ts=[0])


CASES = {"ycsb10k": case_ycsb10k, "mvcc": case_mvcc, "string16": case_string16, "btree_kat": case_btree_kat}


def run_case(case, reader):
    """reader(kind, ...) -> results; returns the flat dict that is stored in / compared with the fixture."""
    out = {}
    for ts in case["probe_ts"]:
        found, pts, key, pay = reader.probe(case["probes"], ts)
        out[f"probe_ts{ts}_found"] = found
        out[f"probe_ts{ts}_stamp"] = pts
        out[f"probe_ts{ts}_crc"] = row_crc(key, pay, found)
    for ts in case["scan_ts"]:
        for S in case["scan_sizes"]:
            for vis in (0, 1):
                counts, key, sts, pay = reader.scan(case["scans"], S, ts, vis)
                valid = scan_valid(counts, S)
                tag = f"scan_ts{ts}_S{S}_v{vis}"
                out[tag + "_counts"] = counts
                out[tag + "_keys"] = np.where(valid[..., None], key, 0).astype(np.uint8)
                out[tag + "_stamp"] = np.where(valid, sts, 0).astype(np.uint32)
                out[tag + "_crc"] = row_crc(key, pay, valid)
    return out
