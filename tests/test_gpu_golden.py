"""GPU: the CUDA path against the committed golden fixtures (outputs of the unmodified reference),
including the reference's own RangeScanBySize known-answer test (test/testing_btree.cpp:630-675)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import cases  # noqa: E402
import stage_indexorganized_b200 as sg  # noqa: E402
from oracle import pyoracle  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_cuda_path_matches_reference_golden(name):
    want = np.load(os.path.join(HERE, "golden", name + ".npz"))
    case = cases.CASES[name](pyoracle.OracleTable)      # the host tree is only the INPUT of the flattening
    with sg.Snapshot.from_tree(case["table"].root, case["key_size"], case["payload"], case["leaf"]) as snap:
        got = cases.run_case(case, snap)
    for k in want.files:
        assert np.array_equal(got[k], want[k]), f"{name}:{k}"
    if name == "btree_kat":
        assert want["scan_ts0_S100_v0_counts"][0] == 100 and want["scan_ts0_S1000_v0_counts"][0] == 1000
