"""Edge shapes of p3_main on the GPU: single level, single column, level counts that are not a multiple of the warp
width, of the 16-byte flag loads or of the phase-A tile, windows of every slot class (1, 2, 4 and, above 128 levels, 8
register slots per lane; the rolled shared-memory window of ice at 97-128 levels), the maximum level count
(P3B200_MAX_NLEV = 256), column counts around the 32-column tile, and the empty grid.

The reference's own size coverage is its unit tests on 1-3 columns x 72 levels and p3_run_and_cmp on ncol x 72
(tests/p3_main_unit_tests.cpp, p3_run_and_cmp.cpp:218-352); sizes are otherwise free parameters of P3Infrastructure
(p3_functions.hpp:127-141), so the kernels have to be right for all of them.  Arithmetic: the parity build against the
oracle on shared transcendentals (see test_gpu_parity_all_columns.py) — every output of every column BIT FOR BIT and
the integer diagnostics equal; the product build on the same shapes must raise no column error and conserve water.
"""
import numpy as np
import pytest

from scream_b200 import conservation, ic_cases
from scream_b200.fields import OUTPUTS
from tests import parity

pytestmark = pytest.mark.gpu

#         ncol  nlev   dt
SHAPES = [(1, 1, 300.0), (40, 1, 1800.0), (1, 128, 300.0), (5, 2, 300.0), (3, 31, 1800.0), (37, 33, 1800.0), (64, 97, 300.0),
          (31, 100, 1800.0), (33, 129, 300.0), (48, 200, 300.0), (40, 255, 300.0), (16, 256, 1800.0),
          (4097, 72, 1800.0)]


@pytest.fixture(scope="module")
def shared_oracle(oracle):
    oracle.set_math_mode("shared")
    yield oracle
    oracle.set_math_mode("libm")


@pytest.fixture(scope="module")
def parity64(p3lib):
    h = p3lib.P3Handle(0, 8, flavour="parity")
    yield h
    h.close()


@pytest.fixture(scope="module")
def parity32(p3lib):
    h = p3lib.P3Handle(0, 4, flavour="parity")
    yield h
    h.close()


def _columns(ncol, nlev, seed, dtype=np.float64):
    if nlev > 1:
        return ic_cases.make_random(ncol, nlev, seed=seed, dtype=dtype)
    # a single level: the lowest level of 16-level columns (make_random spreads its profile over >= 2 levels)
    f = ic_cases.make_random(ncol, 16, seed=seed, dtype=dtype)
    return {n: (np.ascontiguousarray(a[:, 15:]) if a.ndim == 2 else a) for n, a in f.items()}


def _bit_identical(ref, got, d_ref, d_got, what):
    for n in OUTPUTS:
        assert np.array_equal(ref[n], got[n], equal_nan=True), (what, n)
    assert np.array_equal(d_ref[:, :12], d_got[:, :12]), what


@pytest.mark.parametrize("ncol,nlev,dt", SHAPES)
def test_edge_shapes_double(shared_oracle, parity64, ncol, nlev, dt):
    f = _columns(ncol, nlev, 1000 + 7 * nlev + ncol)
    ref, got, d_ref, d_got = parity.run_pair(shared_oracle, parity64, f, dt, True, True)
    _bit_identical(ref, got, d_ref, d_got, (ncol, nlev))
    assert parity64.error_summary() == 0


@pytest.mark.parametrize("ncol,nlev,dt", [(1, 1, 300.0), (37, 33, 1800.0), (64, 97, 300.0), (48, 200, 300.0), (16, 256, 300.0)])
def test_edge_shapes_float(shared_oracle, parity32, ncol, nlev, dt):
    f = _columns(ncol, nlev, 2000 + 7 * nlev + ncol, np.float32)
    ref, got, d_ref, d_got = parity.run_pair(shared_oracle, parity32, f, dt, True, True)
    _bit_identical(ref, got, d_ref, d_got, (ncol, nlev))


def test_edge_shapes_two_steps_tall_columns(shared_oracle, parity64):
    """Two consecutive steps at 200 levels: the second starts from sedimented profiles (8-slot windows, fronts that
    have moved)."""
    f = ic_cases.make_random(24, 200, seed=77)
    ref, got, d_ref, d_got = parity.run_pair(shared_oracle, parity64, f, 300.0, True, False, nsteps=2)
    _bit_identical(ref, got, d_ref, d_got, "24x200 two steps")


@pytest.mark.parametrize("ncol,nlev,dt", [(33, 129, 300.0), (48, 200, 300.0), (16, 256, 1800.0), (64, 97, 300.0)])
def test_edge_shapes_product_build_conserves_water(p3lib, ncol, nlev, dt):
    h = p3lib.P3Handle(0, 8)
    f = ic_cases.make_random(ncol, nlev, seed=3000 + nlev)
    parity.zero_outputs(f)
    dev = p3lib.to_device(f)
    water, energy = conservation.checked_step(h, dev, dt)
    assert abs(water) < 1e-12 and abs(energy) < 1e-12, (water, energy)
    assert h.error_summary() == 0
    h.close()


def test_empty_grid_and_level_limit(p3lib):
    import torch
    h = p3lib.P3Handle(0, 8)
    f = ic_cases.make_random(4, 72)
    dev = p3lib.to_device(f)
    before = {n: v.clone() for n, v in dev.items()}
    # no column: the call succeeds and touches nothing (the struct is built by hand: empty tensors have no address)
    fs, _, _ = h.fields_struct(dev, host=False)
    infra = h._infra(0, 72, 300.0, True, True, True)
    h.main_prepared(fs, infra)
    torch.cuda.synchronize()
    for n in dev:
        assert torch.equal(dev[n].view(torch.int64), before[n].view(torch.int64)), n
    # more levels than the warp-per-column layout holds: rejected, nothing launched
    infra = h._infra(4, 257, 300.0, True, True, True)
    with pytest.raises(p3lib.P3Error):
        h.main_prepared(fs, infra)
    h.close()
