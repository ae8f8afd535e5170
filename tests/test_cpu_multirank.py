"""N > 1 host logic on CPU: world_size-2 gloo processes shard the columns like bench.py does, run the
oracle on their shard, and the one collective of the path (the budget all-reduce) must reproduce the
single-rank totals.  Also pins the oracle's own water/energy conservation."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from scream_b200 import conservation, ic_cases
from tests import parity

NCOL, NLEV, DT = 96, 72, 300.0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tests import oracle_lib
    f = ic_cases.make_random(NCOL, NLEV)
    lo, hi = conservation.shard_columns(NCOL, rank, world)
    shard = {k: v[lo:hi].copy() for k, v in f.items()}
    parity.zero_outputs(shard)
    before = conservation.host_totals(shard, DT)
    oracle_lib.p3_main(shard, DT, True, True, nthreads=1)
    after = conservation.host_totals(shard, DT)
    b, a = conservation.allreduce_totals(before, after)
    if rank == 0:
        out.put((b, a))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_columns_partitions_exactly():
    for n in (1, 7, 96, 1572864):
        for w in (1, 2, 3, 8):
            edges = [conservation.shard_columns(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(300)
def test_two_rank_budget_matches_single_rank(oracle):
    f = ic_cases.make_random(NCOL, NLEV)
    parity.zero_outputs(f)
    before1 = conservation.host_totals(f, DT)
    oracle.p3_main(f, DT, True, True)
    after1 = conservation.host_totals(f, DT)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    b2, a2 = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    np.testing.assert_allclose(b2, before1, rtol=1e-14)
    np.testing.assert_allclose(a2, after1, rtol=1e-13, atol=1e-30)
    water, energy = conservation.budget_residuals(b2, a2)
    # p3_main moves water between species and out through the surface and heats theta with the same constant
    # latent heats the energy measure uses: both budgets close to rounding
    assert abs(water) < 1e-13, water
    assert abs(energy) < 1e-13, energy
