#!/usr/bin/env python
"""Generate the regression fixtures under tests/golden/ from the oracle.

These files pin the oracle (and through it the CUDA path) to ITSELF across commits, integer diagnostics included: any
change of the restatement's results shows up as a diff.  The fixtures that pin the oracle to the REFERENCE are the
ref_*.npz files written by tools/make_ref_golden.py from the reference's own code (oracle/_ref).
Case: ic::Factory::mixed(3, 72), dt = 300, one p3_main call, the four (predictNc, prescribedCCN)
combinations — the reference's p3_run_and_cmp protocol (p3/tests/p3_run_and_cmp.cpp:218-352), first step.

usage: python tools/make_golden.py   (rewrites tests/golden/p3_main_mixed_3x72_dt300.npz)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scream_b200 import ic_cases  # noqa: E402
from scream_b200.fields import OUTPUTS  # noqa: E402
from tests import oracle_lib, parity  # noqa: E402


def main():
    out = {}
    for pn in (0, 1):
        for pc in (0, 1):
            f = ic_cases.make_mixed(3, 72)
            parity.zero_outputs(f)
            _, d = oracle_lib.p3_main(f, 300.0, bool(pn), bool(pc))
            for n in OUTPUTS:
                out[f"nc{pn}_ccn{pc}/{n}"] = f[n]
            out[f"nc{pn}_ccn{pc}/diag"] = d
    path = os.path.join(ROOT, "tests", "golden", "p3_main_mixed_3x72_dt300.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
