"""Generates tests/golden/*.npz by running the recipes of cases.py on the LIVE, UNMODIFIED reference
(oracle/_ref/libstage_ref.so, built from /root/reference by oracle/Makefile.ref).

    python tests/golden/make_golden.py

The fixtures travel with the repository, so the pin on the C oracle and on the CUDA path also holds
where /root/reference does not exist (the GPU box)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

from oracle import pyoracle  # noqa: E402
import cases  # noqa: E402

if __name__ == "__main__":
    pyoracle.build_oracle()
    assert pyoracle.have_ref(), "oracle/_ref/libstage_ref.so missing: make -C oracle -f Makefile.ref"
    for name, fn in cases.CASES.items():
        case = fn(pyoracle.RefTable)
        out = cases.run_case(case, case["table"])
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, len(out), "arrays", os.path.getsize(path), "bytes")
