"""Global water / energy conservation check around p3_main — the ONE collective of the path.

Columns never interact inside p3_main (p3_main_impl.hpp:113), so ranks own disjoint column blocks and the
step itself needs no communication.  What is global is the check that the step conserved water and moist
energy: each rank forms its domain totals on the device (p3b200_water_energy_totals, fixed summation order,
double), and ONE all-reduce of 8 doubles (before + after) over NCCL yields the global budget.  The column
formulas are those of share/property_checks/mass_and_energy_column_conservation_check.cpp:283-363 (which the
reference evaluates per column with a MaxLoc reduction and no MPI, :144-205; the global sum is this build's
addition).  Off the timed path of bench.py.
"""
import numpy as np


def shard_columns(ncol_global, rank, world):
    """Block partition used everywhere (bench.py, tests): rank r owns [lo, hi)."""
    return rank * ncol_global // world, (rank + 1) * ncol_global // world


def budget_residuals(before, after):
    """Relative residuals of the water and energy budgets from two totals vectors
    [water, energy, surface water out, surface energy out] (before: surface terms are zero)."""
    before = np.asarray(before, dtype=np.float64)
    after = np.asarray(after, dtype=np.float64)
    water = (after[0] + after[2] - before[0]) / before[0] if before[0] != 0 else after[0] + after[2]
    energy = (after[1] + after[3] - before[1]) / before[1] if before[1] != 0 else after[1] + after[3]
    return float(water), float(energy)


def allreduce_totals(before, after, group=None, device=None):
    """Sum the two 4-vectors over the ranks of `group` with one all-reduce (NCCL on GPUs, gloo on CPU).
    Returns (before_global, after_global) as numpy arrays; a no-op without an initialised process group."""
    import torch
    import torch.distributed as dist
    buf = torch.tensor(np.concatenate([np.asarray(before, dtype=np.float64), np.asarray(after, dtype=np.float64)]),
                       dtype=torch.float64, device=device if device is not None else "cpu")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    out = buf.cpu().numpy()
    return out[:4].copy(), out[4:].copy()


def host_totals(fields, dt):
    """numpy statement of the same totals, for host-resident fields (tests; oracle outputs)."""
    g, cp, latvap, latice, rhow = 9.80616, 1004.64, 2501000.0, 333700.0, 1000.0
    f = {k: np.asarray(v, dtype=np.float64) for k, v in fields.items()}
    m = f["dpres"] / g
    T = f["th"] / f["inv_exner"]
    water = ((f["qv"] + f["qc"] + f["qi"] + f["qr"]) * m).sum()
    energy = ((T * cp + (latvap + latice) * f["qv"] + latice * (f["qc"] + f["qr"])) * m).sum()
    pw = (rhow * (f["precip_liq_surf"] + f["precip_ice_surf"]) * dt).sum()
    pe = (rhow * latice * f["precip_liq_surf"] * dt).sum()
    return np.array([water, energy, pw, pe])


def checked_step(handle, fields, dt, predict_nc=True, prescribed_ccn=True, group=None):
    """p3_main on this rank's shard bracketed by the budget totals; returns the GLOBAL relative residuals
    (water, energy).  `fields`: dict of device tensors (scream_b200.lib.to_device)."""
    fields["precip_liq_surf"].zero_()
    fields["precip_ice_surf"].zero_()
    before = handle.water_energy_totals(fields, dt)
    handle.main(fields, dt, predict_nc, prescribed_ccn)
    after = handle.water_energy_totals(fields, dt)
    dev = fields["qc"].device
    b, a = allreduce_totals(before, after, group=group, device=dev)
    return budget_residuals(b, a)
