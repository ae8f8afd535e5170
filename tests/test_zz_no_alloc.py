"""p3b200_water_energy_totals_static — the allocation-free variant of the water/energy budget (include/p3_b200.h; no
counterpart in the reference) against the existing entry point
p3b200_water_energy_totals, through the C ABI.

The two run the same three kernels in the same fixed summation order, one with its scratch in the handle's
cudaMalloc'ed block, the other in the library's file-scope __device__ array, so the bar is the one the budget tests use
for repeated evaluation of one shard (tests/test_gpu_conservation.py): the same bits.

The file name sorts last on purpose: a failure here must not hide other tests.
The array is shared by every call and keeps the last call's contents, so the calls below come in a row with shrinking
and growing sizes: stale values of a larger call would show in the smaller one that follows.
"""
import numpy as np
import pytest

from scream_b200 import ic_cases
from tests import parity

pytestmark = pytest.mark.gpu


def _state(ncol, nlev, dtype, seed):
    f = ic_cases.make_random(ncol, nlev, dtype=dtype, seed=seed)
    parity.zero_outputs(f)
    # something in the surface precipitation terms, so that all four totals are exercised
    rng = np.random.default_rng(seed + 1)
    f["precip_liq_surf"] = rng.uniform(0, 1e-6, ncol).astype(dtype)
    f["precip_ice_surf"] = rng.uniform(0, 1e-7, ncol).astype(dtype)
    return f


@pytest.mark.parametrize("rb", [8, 4])
def test_static_scratch_budget_matches_the_allocating_entry_point(p3lib, rb):
    import torch
    dtype = np.float64 if rb == 8 else np.float32
    h = p3lib.P3Handle(0, rb)
    out = torch.full((4,), float("nan"), dtype=torch.float64, device="cuda:0")
    # two calls in a row (and two more): large, small, one column, large again with other data
    cases = [(4096, 72, 11), (37, 128, 12), (1, 72, 13), (p3lib.STATIC_BUDGET_MAX_COLS, 16, 14)]
    devs = [p3lib.to_device(_state(ncol, nlev, dtype, seed)) for ncol, nlev, seed in cases]
    want = [h.water_energy_totals(dev, 300.0) for dev in devs]
    torch.cuda.synchronize()
    free_before = None
    for i, (dev, w) in enumerate(zip(devs, want)):
        out.fill_(float("nan"))
        h.water_energy_totals_static(dev, 300.0, out)
        got = out.cpu().numpy()
        assert np.all(np.isfinite(got)) and got[0] > 0 and got[2] > 0, got
        assert np.array_equal(got, w), (cases[i], got, w)
        # no device memory is taken by a call (the first one may load the module lazily)
        free = torch.cuda.mem_get_info(0)[0]
        if free_before is not None:
            assert free == free_before, (free_before, free)
        free_before = free
    # the same call again gives the same bits, and the allocating entry point still agrees afterwards
    h.water_energy_totals_static(devs[1], 300.0, out)
    assert np.array_equal(out.cpu().numpy(), want[1])
    assert np.array_equal(h.water_energy_totals(devs[1], 300.0), want[1])
    h.close()


def test_static_scratch_budget_on_a_side_stream(p3lib):
    import torch
    h = p3lib.P3Handle(0, 8)
    dev = p3lib.to_device(_state(513, 72, np.float64, 21))
    want = h.water_energy_totals(dev, 300.0)
    out = torch.full((4,), float("nan"), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    h.water_energy_totals_static(dev, 300.0, out, stream=s.cuda_stream)
    s.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)
    h.close()


def test_static_scratch_budget_rejects_more_columns_than_it_holds(p3lib):
    import torch
    h = p3lib.P3Handle(0, 8)
    ncol = p3lib.STATIC_BUDGET_MAX_COLS + 1
    dev = p3lib.to_device(_state(ncol, 8, np.float64, 31))
    out = torch.full((4,), -7.0, dtype=torch.float64, device="cuda:0")
    with pytest.raises(p3lib.P3Error) as err:
        h.water_energy_totals_static(dev, 300.0, out)
    msg = str(err.value)
    assert "rc=-7" in msg and "P3B200_STATIC_BUDGET_MAX_COLS" in msg and str(ncol) in msg, msg
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), np.full(4, -7.0))      # nothing was launched
    # the allocating entry point takes the same columns
    assert np.all(np.isfinite(h.water_energy_totals(dev, 300.0)))
    h.close()
