"""Drop-in parity at the C++ boundary: tests/cpp/dropin_test.cpp runs the UNMODIFIED reference executors
(IndexScanExecutor / TableScanExecutor / BTree::Read / Iterator::GetNext, with a real SSNTransactionManager) and
the header-only adapters of include/stage_gpu_executors.h over the same reference-built tree in one process and
compares every returned struct byte for byte. The binary is built by oracle/Makefile.ref where /root/reference
exists and travels to the GPU box inside oracle/_ref/."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "dropin_test")


def _run(*args):
    return subprocess.run([BIN, *args], capture_output=True, text=True, timeout=600)


@pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/dropin_test not built (no /root/reference here)")
def test_reference_half_runs_on_cpu():
    r = _run("--ref-only", "--rows", "6000", "--probes", "500")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failure(s)" in r.stdout


@pytest.mark.gpu
def test_adapters_match_reference_executors():
    assert os.path.exists(BIN), "oracle/_ref/dropin_test must travel with the snapshot (built by __graft_entry__.build())"
    r = _run()
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "drop-in parity: 0 failure(s)" in r.stdout
    for what in ("ycsb-int", "ycsb-string", "stock"):
        assert any(line.startswith(what) and line.rstrip().endswith("OK") for line in r.stdout.splitlines()), r.stdout
