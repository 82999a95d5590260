"""The drivers' key-stream generators exist twice on purpose — stage_indexorganized_b200/ycsb.py (product side, used by
bench.py's GPU legs) and oracle/pyoracle.py (checker side), so that the product never imports oracle/. They restate the
same code (/root/reference/benchmark/benchmark_common.h:12-105) and must produce the same streams."""
import numpy as np

from oracle import pyoracle
from stage_indexorganized_b200 import ycsb


def test_fastrandom_streams_agree():
    for seed in (1, 2, 1804289383):
        assert np.array_equal(ycsb.lcg_states(seed, 200_000), pyoracle.lcg_states(seed, 200_000))
        assert np.array_equal(ycsb.uniform_keys(10_000, 70_000, seed), pyoracle.uniform_keys(10_000, 70_000, seed))
        assert np.array_equal(ycsb.random_strings(seed, 5000, 16), pyoracle.random_strings(seed, 5000, 16))
    fr = pyoracle.FastRandom(9)
    assert [int(x) for x in ycsb.fastrandom_next64(9, 6)] == [fr.next() for _ in range(6)]


def test_zipf_streams_agree():
    for n, theta in ((9999, 0.2), (99_999, 0.9)):
        assert np.array_equal(ycsb.Zipf(n, theta, seed=1).draw(60_000), pyoracle.Zipf(n, theta, seed=1).draw(60_000))
