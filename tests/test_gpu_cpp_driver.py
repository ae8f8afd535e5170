"""The C++ caller written against the reference-shaped face p3::Functions<Real,B200Device>
(tests/cpp/p3_run_and_cmp_b200.cpp, the analogue of p3/tests/p3_run_and_cmp.cpp) vs the oracle."""
import os
import subprocess

import numpy as np
import pytest

from scream_b200 import ic_cases, lib as p3
from scream_b200.fields import FIELDS, OUTPUTS, levels_of
from tests import parity

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "p3_run_and_cmp_b200")


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "cpp"), "-s"])


def test_cpp_driver_compiles_and_links():
    """CPU: the reference-shaped header compiles as C++17 and links against libp3b200.so."""
    _build()
    assert os.path.exists(EXE)
    # without arguments it prints its usage and returns 1 (no CUDA call is made)
    assert subprocess.run([EXE], capture_output=True).returncode == 1


@pytest.mark.gpu
@pytest.mark.parametrize("predict_nc,prescribed_ccn", [(False, False), (True, True)])
def test_cpp_driver_six_steps(oracle, tmp_path, predict_nc, prescribed_ccn):
    """p3_run_and_cmp protocol: ic::Factory::mixed(3,72), dt=300, 6 steps, every output field compared."""
    _build()
    ncol, nlev, dt, nsteps = 3, 72, 300.0, 6
    f = ic_cases.make_mixed(ncol, nlev)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as fh:
        for n in FIELDS:
            fh.write(np.ascontiguousarray(f[n], dtype=np.float64).tobytes())
    subprocess.check_call([EXE, p3.TABLE_BLOB, str(fin), str(fout), str(ncol), str(nlev), str(dt), str(nsteps),
                           str(int(predict_nc)), str(int(prescribed_ccn))])
    raw = np.fromfile(fout, dtype=np.float64)
    got, off = {}, 0
    for n in FIELDS:
        lev = levels_of(n, nlev)
        got[n] = raw[off:off + ncol * lev].reshape((ncol, lev) if f[n].ndim == 2 else (ncol,))
        off += ncol * lev
    ref = parity.clone(f)
    for _ in range(nsteps):
        parity.zero_outputs(ref)
        oracle.p3_main(ref, dt, predict_nc, prescribed_ccn)
    errs = parity.field_errors(ref, got, names=OUTPUTS)
    assert max(errs.values()) <= 1e-9, {k: v for k, v in errs.items() if v > 1e-9}
