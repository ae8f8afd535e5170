"""Oracle vs the committed regression fixtures (tests/golden/, written by tools/make_golden.py), and the
GPU path vs the same fixtures (so the GPU box needs neither /root/reference nor a fresh oracle run for it)."""
import os

import numpy as np
import pytest

from scream_b200 import ic_cases
from scream_b200.fields import OUTPUTS
from tests import parity

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "p3_main_mixed_3x72_dt300.npz")
FLAGS = [(0, 0), (0, 1), (1, 0), (1, 1)]


@pytest.mark.parametrize("pn,pc", FLAGS)
def test_oracle_reproduces_golden(oracle, pn, pc):
    g = np.load(GOLDEN)
    f = ic_cases.make_mixed(3, 72)
    parity.zero_outputs(f)
    _, d = oracle.p3_main(f, 300.0, bool(pn), bool(pc))
    assert np.array_equal(d, g[f"nc{pn}_ccn{pc}/diag"])
    for n in OUTPUTS:
        # same compiler flags -> bit-identical; allow a last-bit libm difference between glibc versions
        np.testing.assert_allclose(f[n], g[f"nc{pn}_ccn{pc}/{n}"], rtol=1e-13, atol=1e-300, err_msg=n)


@pytest.mark.gpu
@pytest.mark.parametrize("pn,pc", FLAGS)
def test_gpu_reproduces_golden(p3lib, pn, pc):
    g = np.load(GOLDEN)
    h = p3lib.P3Handle(0, 8)
    f = ic_cases.make_mixed(3, 72)
    parity.zero_outputs(f)
    dev = p3lib.to_device(f)
    h.main(dev, 300.0, bool(pn), bool(pc))
    got = p3lib.to_host(dev)
    ref = {n: g[f"nc{pn}_ccn{pc}/{n}"] for n in OUTPUTS}
    errs = parity.field_errors(ref, got)
    assert max(errs.values()) <= parity.TOL[8], {k: v for k, v in errs.items() if v > parity.TOL[8]}
    assert np.array_equal(h.get_diag(3), g[f"nc{pn}_ccn{pc}/diag"])
    h.close()
