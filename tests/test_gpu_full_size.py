"""Size-independent properties of the CUDA p3_main at a BASELINE.json grid size (ne120pg2: 345 600 columns x 128
levels, dt = 300 s, fp64) — the oracle cannot run that many columns in seconds, so the full-size run is checked by

  * replica invariance: the grid tiles 96 distinct columns; every replica of a column must be BIT-identical to the
    first one in all 35 outputs and all integer diagnostics (any dependence on the tile/work-queue/sort order, a
    race on the shared accumulators or a stale scratch value would break this);
  * the first 96 columns against the CPU oracle (tolerance 1e-12 on the well-conditioned columns, integer
    diagnostics bit-exact);
  * the global water budget closing to 1e-12 (relative) and no column error bit.
"""
import os

import numpy as np
import pytest

from scream_b200 import conservation, ic_cases
from scream_b200.fields import FIELDS, OUTPUTS
from tests import parity

pytestmark = pytest.mark.gpu

# P3B200_TEST_FULL_NCOL=1572864 runs the same checks on the ne256pg2 grid (60 GB of fields)
NCOL, NLEV, DT, TILE = int(os.environ.get("P3B200_TEST_FULL_NCOL", 345600)), 128, 300.0, 96


def _device_grid(torch, tile, ncol):
    dev = torch.device("cuda", 0)
    idx = torch.arange(ncol, device=dev) % TILE
    return {n: torch.from_numpy(tile[n]).to(dev).index_select(0, idx).contiguous() for n in FIELDS}


@pytest.mark.parametrize("ic", ["mixed", "random"])
def test_full_size_replicas_budget_and_oracle(oracle, p3lib, ic):
    torch = pytest.importorskip("torch")
    if ic == "mixed":
        tile = ic_cases.make_mixed_tiled(TILE, NLEV, tile=3, dtype=np.float64)
    else:
        tile = ic_cases.make_random(TILE, NLEV, dtype=np.float64)
    parity.zero_outputs(tile)
    h = p3lib.P3Handle(0, 8)
    fields = _device_grid(torch, tile, NCOL)
    before = h.water_energy_totals(fields, DT)
    h.main(fields, DT, True, True)
    after = h.water_energy_totals(fields, DT)
    assert h.error_summary() == 0
    water, energy = conservation.budget_residuals(before, after)
    assert abs(water) < 1e-12, water

    # replica invariance, bit for bit
    for n in OUTPUTS:
        t = fields[n]
        first = t[:TILE]
        rep = t.view(NCOL // TILE, TILE, *t.shape[1:])
        same = (rep.view(torch.int64) == first.view(torch.int64).unsqueeze(0)).all()
        assert bool(same), f"{n}: replicas of a column differ"
    diag = h.get_diag(NCOL)
    assert np.array_equal(diag.reshape(NCOL // TILE, TILE, -1), np.broadcast_to(diag[:TILE], (NCOL // TILE, TILE, diag.shape[1])))

    # the first tile against the oracle
    got = {n: fields[n][:TILE].cpu().numpy() for n in FIELDS}
    ref = {n: v.copy() for n, v in tile.items()}
    _, d_ref = oracle.p3_main(ref, DT, True, True, want_diag=True)
    stable = parity.stable_columns(oracle, tile, DT, True, True)
    assert stable.sum() >= TILE // 2
    errs = parity.field_errors({n: ref[n][stable] for n in OUTPUTS}, {n: got[n][stable] for n in OUTPUTS})
    assert max(errs.values()) <= parity.TOL[8], {k: v for k, v in errs.items() if v > parity.TOL[8]}
    assert np.array_equal(d_ref[stable], diag[:TILE][stable])
    h.close()
