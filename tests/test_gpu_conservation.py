"""The device budget kernels (p3b200_water_energy_totals) and the conservation of the CUDA p3_main."""
import numpy as np
import pytest

from scream_b200 import conservation, ic_cases
from tests import parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rb,tol", [(8, 1e-13), (4, 1e-5)])
def test_device_totals_and_conservation(p3lib, rb, tol):
    dtype = np.float64 if rb == 8 else np.float32
    h = p3lib.P3Handle(0, rb)
    f = ic_cases.make_random(5000, 128, dtype=dtype)
    parity.zero_outputs(f)
    dev = p3lib.to_device(f)
    before = h.water_energy_totals(dev, 300.0)
    np.testing.assert_allclose(before, conservation.host_totals(f, 300.0), rtol=1e-13, atol=1e-30)
    water, energy = conservation.checked_step(h, dev, 300.0)
    assert abs(water) < tol and abs(energy) < tol, (water, energy)
    after = h.water_energy_totals(dev, 300.0)
    np.testing.assert_allclose(after, conservation.host_totals(p3lib.to_host(dev), 300.0), rtol=1e-13, atol=1e-30)
    # fixed summation order: the same shard gives the same bits
    assert np.array_equal(after, h.water_energy_totals(dev, 300.0))
    assert after[2] > 0            # something precipitated
    h.close()
