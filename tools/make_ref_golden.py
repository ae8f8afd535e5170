#!/usr/bin/env python
"""Generate tests/golden/ref_*.npz from the REFERENCE'S OWN p3_main (oracle/_ref, built by oracle/Makefile.ref from the
unmodified sources under /root/reference against the Kokkos stand-in oracle/ref_shim/).  Runs only in the container
that has /root/reference; the fixtures are committed so that every other machine (the GPU box included) can check the
oracle and the CUDA path against outputs of the reference itself.

Cases (protocol of the reference's p3_run_and_cmp, p3/tests/p3_run_and_cmp.cpp:218-352):
  ref_mixed_3x72_dt300_6steps.npz   ic::Factory::mixed(3, 72), dt = 300, SIX consecutive steps, the four
                                    (predictNc, prescribedCCN) combinations, fp64 and fp32, math = libm and shared
  ref_random_{72x1800,128x300}.npz  64 seeded random-activity columns (G2, seed 20241104), one step, flags (T,T),
                                    fp64 and fp32, math = libm and shared
"math = shared" is the hooked reference library running on the deterministic transcendentals of
scream_b200/csrc/p3_math.hpp — the arithmetic of the GPU parity build.

usage: python tools/make_ref_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scream_b200 import ic_cases  # noqa: E402
from scream_b200.fields import OUTPUTS  # noqa: E402
from tests import parity, ref_lib  # noqa: E402

NRAND = 64
SEED = 20241104


def as_dtype(f, rb):
    return {k: v.astype(np.float32) for k, v in f.items()} if rb == 4 else f


def main():
    assert ref_lib.can_build(), "needs /root/reference"
    ref_lib.build()
    out = {}
    for rb in (8, 4):
        for math in ("libm", "shared"):
            for pn in (0, 1):
                for pc in (0, 1):
                    f = as_dtype(ic_cases.make_mixed(3, 72), rb)
                    for step in range(6):
                        parity.zero_outputs(f)
                        ref_lib.p3_main(f, 300.0, bool(pn), bool(pc), math=math)
                        for n in OUTPUTS:
                            out[f"f{rb * 8}/{math}/nc{pn}_ccn{pc}/step{step}/{n}"] = f[n].copy()
    path = os.path.join(ROOT, "tests", "golden", "ref_mixed_3x72_dt300_6steps.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(out), "arrays")

    for nlev, dt in ((72, 1800.0), (128, 300.0)):
        out = {}
        for rb in (8, 4):
            for math in ("libm", "shared"):
                f = as_dtype(ic_cases.make_random(NRAND, nlev, seed=SEED), rb)
                parity.zero_outputs(f)
                ref_lib.p3_main(f, dt, True, True, math=math)
                for n in OUTPUTS:
                    out[f"f{rb * 8}/{math}/{n}"] = f[n].copy()
        path = os.path.join(ROOT, "tests", "golden", f"ref_random_{nlev}x{int(dt)}.npz")
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
