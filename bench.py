#!/usr/bin/env python
"""bench.py — P3 p3_main throughput (column-levels/s) on B200, with roofline, e2e and CPU baseline.

Contract: `python bench.py --gpus N --steps K --warmup W` (under torchrun for N>1) prints ONE JSON line
on rank 0.  A "step" is one p3_main call over this rank's column shard on a fresh copy of the inputs
(P3 mutates its state; the restore is outside the event-timed region and doubles as the L2 flush).

  value     device-timed, inputs resident in HBM: ncol*nlev*K / sum of per-step CUDA-event times (max over ranks)
  e2e       the same metric through p3b200_main_host with PINNED HOST buffers (H2D + kernel + D2H in the timed region)
  roofline  algorithmic bytes (336 B per column-level in fp64 with default flags) / kernel time
  cpu_baseline  the CPU oracle ("port": restatement of the reference, the reference itself cannot be built here)

`--impl reference` times that CPU oracle on all host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from scream_b200 import ic_cases  # noqa: E402
from scream_b200.fields import (FIELDS, OUTPUTS, PROGNOSTIC, algorithmic_bytes_per_col_level)  # noqa: E402

WORKLOADS = {
    # name: (ncol, nlev, dt, real_bytes)
    "ne30pg2_L72": (21600, 72, 1800.0, 8),
    "ne120pg2_L128": (345600, 128, 300.0, 8),
    "ne256pg2_L128": (1572864, 128, 300.0, 8),
    "ne1024pg2_L128_sp_pergpu": (3145728, 128, 300.0, 4),
}
METRIC = "p3_main_column_levels_per_sec"
UNIT = "column-levels/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="ne256pg2_L128", choices=sorted(WORKLOADS))
    ap.add_argument("--ic", default="mixed", choices=["mixed", "mixed_sp", "random"],
                    help="mixed: tiled ic::Factory::mixed (G1); mixed_sp: the same without its cloudy level at 191 K, where the "
                         "reference itself overflows in single precision; random: seeded per-column cloud layers (G2)")
    ap.add_argument("--ncol", type=int, default=0, help="override the workload's global column count")
    ap.add_argument("--e2e-cols", type=int, default=98304, help="columns of the e2e (host-buffer) sample per rank")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="target CPU work for the cpu_baseline sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-prepost", action="store_true", help="skip the p3_preamble/p3_postamble bandwidth measurement")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--flavour", default="product", choices=["product", "parity"],
                    help="library build: product (ships) or parity (-DP3B200_PARITY, bit-identical to the shared-math oracle)")
    ap.add_argument("--zero-history", type=int, default=0, choices=[0, 1],
                    help="P3B200Infra::zero_history: 1 folds run_impl's zero-fill of the three history arrays into the call")
    ap.add_argument("--no-alt", action="store_true", help="skip the second workload line (random-activity IC, G2)")
    ap.add_argument("--alt-steps", type=int, default=5)
    return ap.parse_args()


def host_threads():
    """Threads the CPU arm may use: the process's CPU affinity, NOT omp_get_max_threads() — torchrun exports
    OMP_NUM_THREADS=1 to its workers, which throttled the round-1 reference arm to one core at N >= 2."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


CPU_SIMD = "scalar restatement (pack size 1), g++ -O3 -march=x86-64-v3, OpenMP over columns"


def workload_config(args, world, ncol_g, ncol, nlev, dt):
    """The `config` object: identical for the B200 arm and the --impl reference arm (which times a bounded sample of it)."""
    return {"workload": args.workload, "ic": args.ic, "global_columns": ncol_g, "nlev": nlev, "dt": dt,
            "columns_per_gpu": ncol, "flags": "predictNc=1,prescribedCCN=1",
            "parallelism": f"columns block-sharded over {world} GPU(s), no collective in the step",
            "l2": "inputs 72 GB >> 126 MB L2 and restored (16 GB written) between timed steps"}


def make_tile(ic, ncols, nlev, dtype):
    """Host tile of distinct columns that the device grid repeats."""
    if ic in ("mixed", "mixed_sp"):
        return ic_cases.make_mixed_tiled(ncols, nlev, tile=3, dtype=dtype, single_precision_safe=(ic == "mixed_sp"))
    return ic_cases.make_random(ncols, nlev, dtype=dtype)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            p = [x.strip() for x in line.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1])); power.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arms.  "reference" = oracle/_ref: the reference's OWN p3_main (unmodified sources compiled against the Kokkos
# stand-in, oracle/Makefile.ref), at the reference's CPU pack size 16 or at pack size 1, whichever is faster on this
# host; "port" = the scalar restatement oracle/ (bit-identical to the reference, tests/test_cpu_oracle_vs_ref.py).
def cpu_arms(real_bytes):
    from tests import oracle_lib, ref_lib
    arms = []
    for pack in (16, 1):
        if ref_lib.available(real_bytes, pack):
            arms.append({"kind": "reference", "pack": pack,
                         "simd": f"reference p3_main, Pack<Real,{pack}>, g++ -O2 -march=x86-64-v3 -fopenmp, serial Kokkos stand-in "
                                 "(one team per column, OpenMP over columns)",
                         "run": (lambda f, dt, n, pack=pack: ref_lib.p3_main(f, dt, True, True, nthreads=n, pack=pack))})
    arms.append({"kind": "port", "pack": 1, "simd": CPU_SIMD,
                 "run": (lambda f, dt, n: oracle_lib.p3_main(f, dt, True, True, want_diag=False, nthreads=n)[0])})
    return arms


def pick_cpu_arm(ic, nlev, dt, real_bytes, threads, prefer_reference):
    """Probe every available CPU arm on a small tile; returns (arm, columns/s of each arm)."""
    dtype = np.float64 if real_bytes == 8 else np.float32
    probe_cols = 96 * max(threads, 1)
    rates = {}
    arms = cpu_arms(real_bytes)
    for arm in arms:
        f = make_tile(ic, probe_cols, nlev, dtype)
        arm["run"](f, dt, threads)                      # first touch / table setup
        f = make_tile(ic, probe_cols, nlev, dtype)
        t = arm["run"](f, dt, threads)
        arm["rate"] = probe_cols / max(t, 1e-9)
        rates[f"{arm['kind']}_pk{arm['pack']}"] = arm["rate"] * nlev
    pool = [a for a in arms if a["kind"] == "reference"] if prefer_reference else arms
    if not pool:
        pool = arms
    return max(pool, key=lambda a: a["rate"]), rates


def cpu_oracle_rate(ic, nlev, dt, real_bytes, target_seconds, predict_nc=True, prescribed_ccn=True):
    """cpu_baseline: the fastest CPU arm (all host threads) on a bounded sample of the workload."""
    dtype = np.float64 if real_bytes == 8 else np.float32
    threads = host_threads()
    arm, rates = pick_cpu_arm(ic, nlev, dt, real_bytes, threads, prefer_reference=False)
    # size the sample for `target_seconds` of CPU WORK (core-seconds), i.e. target/threads of wall
    ncol = int(min(max(arm["rate"] * target_seconds / max(threads, 1), 96 * threads), 400000))
    f = make_tile(ic, ncol, nlev, dtype)
    t = arm["run"](f, dt, threads)
    return {"value": ncol * nlev / t, "unit": UNIT, "cores": threads, "kind": arm["kind"], "simd": arm["simd"],
            "sample": f"{ncol} columns x {nlev} levels of the '{ic}' IC, dt={dt}, one p3_main call, "
                      f"{t:.3f} s on {threads} OpenMP threads",
            "probe_column_levels_per_sec": {k: round(v) for k, v in rates.items()}}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation (oracle/_ref; the oracle port where it is absent)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ncol_g, nlev, dt, rb = WORKLOADS[args.workload]
    dtype = np.float64 if rb == 8 else np.float32
    threads = host_threads()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    arm, rates = pick_cpu_arm(args.ic, nlev, dt, rb, threads, prefer_reference=True)
    # each step ~2 s of wall so that steps+warmup stay within a few minutes
    ncol = int(min(max(arm["rate"] * 2.0, 64 * threads), 400000))
    base = make_tile(args.ic, ncol, nlev, dtype)
    times = []
    for s in range(args.warmup + args.steps):
        f = {k: v.copy() for k, v in base.items()}
        t = arm["run"](f, dt, threads)
        if s >= args.warmup:
            times.append(t)
    total = sum(times)
    value = ncol * nlev * len(times) / total
    what = ("the reference's own p3_main (oracle/_ref: unmodified sources, Kokkos stand-in)" if arm["kind"] == "reference"
            else "CPU restatement (oracle port) of the reference's p3_main; oracle/_ref is not present on this machine")
    sample = (f"{ncol} columns x {nlev} levels of the '{args.ic}' IC per step (a bounded sample of the configured grid), dt={dt}, "
              f"{threads} OpenMP threads; {what}")
    if args.ncol:
        ncol_g = args.ncol
    per_gpu = ncol_g if args.scaling == "weak" else (ncol_g // world)
    glob = ncol_g * world if args.scaling == "weak" else ncol_g
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f64" if rb == 8 else "f32", "data": "synthetic",
            "config": workload_config(args, world, glob, per_gpu, nlev, dt),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": arm["kind"], "simd": arm["simd"],
                             "sample": sample, "sample_columns_per_step": ncol,
                             "probe_column_levels_per_sec": {k: round(v) for k, v in rates.items()}},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from scream_b200 import lib as p3

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ncol_g, nlev, dt, rb = WORKLOADS[args.workload]
    if args.ncol:
        ncol_g = args.ncol
    if args.scaling == "strong":
        lo, hi = rank * ncol_g // world, (rank + 1) * ncol_g // world
        ncol = hi - lo
    else:
        ncol = ncol_g
        ncol_g = ncol * world
    dtype = np.float64 if rb == 8 else np.float32
    tdtype = torch.float64 if rb == 8 else torch.float32
    predict_nc, prescribed_ccn = True, True

    # ---- device-resident inputs: a host tile of distinct columns repeated on the device ----------
    tile_cols = 3 if args.ic in ("mixed", "mixed_sp") else 8192
    tile = make_tile(args.ic, tile_cols, nlev, dtype)
    idx = (torch.arange(ncol, device=dev) + (rank * ncol_g // world if args.scaling == "strong" else 0)) % tile_cols
    fields = {}
    for n in FIELDS:
        t = torch.from_numpy(tile[n]).to(dev)
        fields[n] = t.index_select(0, idx).contiguous()
    pristine = {n: fields[n].clone() for n in PROGNOSTIC}
    zero_names = [n for n in OUTPUTS if n not in PROGNOSTIC]

    def restore():
        for n in PROGNOSTIC:
            fields[n].copy_(pristine[n])
        for n in zero_names:          # run_impl zero-fills the history arrays; p3_main_f all outputs
            fields[n].zero_()

    handle = p3.P3Handle(local, rb, flavour=args.flavour)
    fstruct, _, _ = handle.fields_struct(fields, host=False)
    infra = handle._infra(ncol, nlev, dt, predict_nc, prescribed_ccn, True, zero_history=bool(args.zero_history))
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        restore()
        handle.main_prepared(fstruct, infra, stream)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.time()
    for s in range(args.steps):
        restore()                      # outside the event bracket: fresh inputs + evicts L2 (16 GB >> 126 MB)
        ev[s][0].record()
        handle.main_prepared(fstruct, infra, stream)
        ev[s][1].record()
    barrier()
    t_wall1 = time.time()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = float(sum(step_ms))
    err_bits = handle.error_summary()
    phases = handle.last_phase_ms()      # CUDA-event split of the LAST timed step over the five launches
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    tmax = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms_max = float(tmax.item())
    value = ncol_g * nlev * args.steps / (total_ms_max * 1e-3)

    # ---- the one collective of the path, off the timed region: global water/energy budget of one step
    from scream_b200 import conservation
    restore()
    fields["precip_liq_surf"].zero_(); fields["precip_ice_surf"].zero_()
    b4 = handle.water_energy_totals(fields, dt, stream)
    handle.main_prepared(fstruct, infra, stream)
    a4 = handle.water_energy_totals(fields, dt, stream)
    b4, a4 = conservation.allreduce_totals(b4, a4, device=dev)      # NCCL all-reduce of 8 doubles when world > 1
    res_w, res_e = conservation.budget_residuals(b4, a4)
    budget = {"water_rel_residual": res_w, "energy_rel_residual": res_e, "global_water_kg_m2_sum": float(b4[0]),
              "collective": "all_reduce(8 x f64, NCCL)" if world > 1 else "none (1 rank)"}

    # ---- roofline for the one kernel (p3_main_kernel): algorithmic bytes / event time --------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak = float(json.load(open(peaks_path))["hbm_gbs"]); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak = 6650.0; peak_src = "fallback (B200_PROFILING.md 6.65 TB/s)"
    bpl = algorithmic_bytes_per_col_level(rb, predict_nc, prescribed_ccn)
    n_sed = 3 * (1 if nlev <= 32 else 2 if nlev <= 64 else 3 if nlev <= 128 else 4)   # one launch per species and register-window class
    n_launch = 8 + n_sed + 1     # phase A: classification pass + six part2 groups + column finish
    achieved = bpl * ncol * nlev / (total_ms / args.steps * 1e-3) / 1e9
    # measured DRAM traffic of one p3_main (all launches), from the committed ncu capture of ONE step at the full
    # ne256pg2 grid (profiles/r02o_dram_traffic_full_grid.txt): bytes per column-level of the same IC, scaled to this shard
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            per_cl = json.load(open(tpath)).get(f"{args.ic}_L{nlev}_f{8 * rb}", {}).get("dram_bytes_per_column_level")
            traffic = per_cl * ncol * nlev if per_cl else None
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": f"p3_main = p3_classify + 6 x p3_part2<group> + p3_column_finish + {n_sed} x p3_sediment<cloud|rain|ice, window class> + p3_phase_c ({n_launch} launches, one event bracket); longest single kernel: p3_sediment<rain, 1 slot> (phase_ms_last_step)",
                "bytes_per_column_level": bpl, "peak_source": peak_src}

    # ---- e2e: pinned host buffers through p3b200_main_host ---------------------------------------
    e2e = None
    if not args.no_e2e:
        ecols = min(args.e2e_cols, ncol)
        host = {}
        for n in FIELDS:
            src = fields[n][:ecols] if n not in PROGNOSTIC else pristine[n][:ecols]
            h = torch.empty(src.shape, dtype=tdtype, pin_memory=True)
            h.copy_(src)
            host[n] = h
        host_np = {n: host[n].numpy() for n in FIELDS}
        host0 = {n: host[n].clone() for n in PROGNOSTIC}
        n_in = 10 + 9 + 1          # prognostic + always-read inputs + nccn; the history arrays are zero-filled by the call (zero_history)
        n_out = 22
        h2d = n_in * ecols * nlev * rb
        d2h = n_out * ecols * nlev * rb + 2 * ecols * rb + 2 * ecols * rb  # fluxes have one more level
        times = []
        for s in range(2 + 3):
            for n in PROGNOSTIC:
                host[n].copy_(host0[n])
            for n in zero_names:
                host[n].zero_()
            barrier()
            t0 = time.perf_counter()
            handle.main_host(host_np, dt, predict_nc, prescribed_ccn, want_diag=False, zero_history=True)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            if s >= 2:
                times.append(t1 - t0)
        te = torch.tensor([float(np.mean(times))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": ecols * world * nlev / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "sample_columns_per_gpu": ecols, "ms_per_step": 1e3 * float(te.item()),
               "api": "p3b200_main_host (pinned host arrays, chunked H2D/kernel/D2H over 4 streams, zero_history: the three history arrays are zero-filled on the device instead of copied in)"}

    # ---- the brackets of P3Microphysics::run_impl around p3_main: two pure HBM streams ----
    prepost = None
    if not args.no_prepost:
        try:
            from scream_b200 import prepost_cases
            from scream_b200.fields import PREAMBLE_ARRAYS_RW, POSTAMBLE_ARRAYS_RW
            pcols = min(ncol, 524288)
            t64 = prepost_cases.make_preamble(64, nlev, dtype)
            pidx = torch.arange(pcols, device=dev) % 64
            pre = {n: torch.from_numpy(v).to(dev).index_select(0, pidx).contiguous() for n, v in t64.items()}
            p64 = prepost_cases.make_postamble(64, nlev, dtype)
            post = {n: torch.from_numpy(v).to(dev).index_select(0, pidx).contiguous() for n, v in p64.items()}
            res = {}
            # the fused property checks (§8f N2) are timed on the same arrays: radii zeroed so that repeating the
            # in-place postamble keeps every field inside its interval (no failure-location passes in the timing)
            from scream_b200.fields import CHECK_FIELDS, CHECK_POSTAMBLE_SOURCE
            chk_in = {n: post[CHECK_POSTAMBLE_SOURCE.get(n, n)] for n in CHECK_FIELDS}
            last = {}

            def checked():
                last["r"] = handle.postamble_checked(post, dt)

            def standalone():
                last["r"] = handle.property_checks(chk_in)

            for name, fn, rw in (("preamble", lambda: handle.preamble(pre), PREAMBLE_ARRAYS_RW),
                                 ("postamble", lambda: handle.postamble(post, dt), POSTAMBLE_ARRAYS_RW),
                                 ("postamble_with_15_checks", checked, POSTAMBLE_ARRAYS_RW),
                                 ("property_checks_alone", standalone, (13, 0))):
                if name == "postamble_with_15_checks":
                    for n in ("diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr"):
                        post[n].zero_()
                for _ in range(3):
                    fn()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(5):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 5
                gbs = sum(rw) * rb * pcols * nlev / (ms * 1e-3) / 1e9
                res[name] = {"ms": ms, "columns": pcols, "arrays_read_written": list(rw), "achieved_gbs": gbs,
                             "frac_of_hbm_peak": gbs / peak}
                if "r" in last:
                    res[name]["checks_passed"] = sum(1 for v in last.pop("r").values() if v["result"] == 0)
            prepost = res
            del pre, post
        except Exception as exc:      # never let the side measurement take the headline down
            prepost = {"error": repr(exc)}

    # ---- p3_main + p3_postamble: two calls vs the postamble fused into the tail of phase C, on the bench grid ----
    if prepost is not None and "error" not in prepost:
        try:
            post = {"pmid": fields["pres"], "pseudo_density": fields["dpres"], "th_atm": fields["th"],
                    "pseudo_density_dry": (fields["dpres"] * (1 - fields["qv"])).contiguous(),
                    "T_atm": (fields["th"] / fields["inv_exner"]).contiguous(),
                    "T_prev": torch.empty_like(fields["th"]), "qv_prev": torch.empty_like(fields["th"]),
                    "precip_liq_surf_flux": fields["precip_liq_surf"], "precip_ice_surf_flux": fields["precip_ice_surf"],
                    "precip_liq_surf_mass": torch.zeros_like(fields["precip_liq_surf"]),
                    "precip_ice_surf_mass": torch.zeros_like(fields["precip_ice_surf"])}
            for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "diag_eff_radius_qc", "diag_eff_radius_qi",
                      "diag_eff_radius_qr"):
                post[n] = fields[n]
            t_atm0 = post["T_atm"].clone()

            def two_calls():
                handle.main_prepared(fstruct, infra, stream)
                handle.postamble(post, dt)

            def fused():
                handle.main_postamble(fields, post, dt, predict_nc, prescribed_ccn)

            tail = {}
            for name, fn in (("main_then_postamble", two_calls), ("main_postamble_one_call", fused)):
                ms = []
                for it in range(4):
                    restore()
                    post["T_atm"].copy_(t_atm0)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    fn()
                    e1.record()
                    torch.cuda.synchronize()
                    if it:                      # first one is the warm-up
                        ms.append(e0.elapsed_time(e1))
                tail[name + "_ms"] = sum(ms) / len(ms)
                tail[name + "_phase_c_ms"] = handle.last_phase_ms()["phase_c"]
            # ---- one whole run_impl (eamxx_p3_run.cpp:5-46) in ONE event bracket: p3_preamble, the zero-fill of the three
            # history arrays (folded into p3_main: zero_history), p3_main, p3_postamble with the 15 property checks.
            # Algorithmic bytes per column-level: 336 (p3_main) + 248 (preamble: 15 reads,
            # 16 writes) + 256 (postamble: 17 reads, 15 writes) + 24 (history zero-fill) = 864 in fp64.
            pre_full = {"pmid": post["pmid"], "pseudo_density": post["pseudo_density"], "pseudo_density_dry": post["pseudo_density_dry"],
                        "T_atm": post["T_atm"], "cld_frac_t": torch.maximum(fields["cld_frac_l"], fields["cld_frac_i"]).contiguous(),
                        "qv_prev": fields["qv_prev"], "inv_exner": fields["inv_exner"], "th_atm": fields["th"],
                        "cld_frac_l": fields["cld_frac_l"], "cld_frac_i": fields["cld_frac_i"], "cld_frac_r": fields["cld_frac_r"],
                        "dz": fields["dz"]}
            for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
                pre_full[n] = fields[n]
            infra_zh = handle._infra(ncol, nlev, dt, predict_nc, prescribed_ccn, True, zero_history=True)
            keep = {n: fields[n].clone() for n in ("inv_exner", "cld_frac_l", "cld_frac_i", "cld_frac_r", "dz", "qv_prev")}
            for fused_pre in (False, True):
                ms, parts = [], None
                for it in range(4):
                    restore()
                    post["T_atm"].copy_(t_atm0)
                    for n, v in keep.items():
                        fields[n].copy_(v)
                    ev4 = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                    ev4[0].record()
                    if fused_pre:
                        ev4[1].record()
                        handle.main_preamble(fields, pre_full, dt, predict_nc, prescribed_ccn, zero_history=True)
                    else:
                        handle.preamble(pre_full)
                        ev4[1].record()
                        handle.main_prepared(fstruct, infra_zh, stream)
                    ev4[2].record()
                    if fused_pre:
                        tail["run_impl_p3_main_phases_ms"] = {k: round(v, 3) for k, v in handle.last_phase_ms().items()}
                    chk = handle.postamble_checked(post, dt)
                    ev4[3].record()
                    torch.cuda.synchronize()
                    if it:
                        ms.append(ev4[0].elapsed_time(ev4[3]))
                        parts = [ev4[i].elapsed_time(ev4[i + 1]) for i in range(3)]
                run_ms = sum(ms) / len(ms)
                bytes_cl = (42 + 31 + 32 + 3) * rb
                gbs = bytes_cl * ncol * nlev / (run_ms * 1e-3) / 1e9
                if not fused_pre:
                    tail["run_impl_three_calls_ms"] = run_ms
                    tail["run_impl_three_calls_parts_ms"] = {"preamble": parts[0], "p3_main_zero_history": parts[1],
                                                             "postamble_with_15_checks": parts[2]}
                    continue
                # the headline of the bracket: p3_preamble fused into p3_main's classification pass (p3b200_main_preamble)
                tail["run_impl_ms"] = run_ms
                tail["run_impl_parts_ms"] = {"p3_main_with_fused_preamble_and_zero_history": parts[1], "postamble_with_15_checks": parts[2]}
                tail["run_impl_roofline"] = {"bound": "hbm", "bytes_per_column_level": bytes_cl, "achieved": gbs, "peak": peak,
                                             "unit": "GB/s", "frac": gbs / peak}
                tail["run_impl_checks_passed"] = sum(1 for v in chk.values() if v["result"] == 0)
            for n, v in keep.items():
                fields[n].copy_(v)
            del pre_full, keep
            tail["columns"] = ncol
            prepost["run_impl_tail"] = tail
            del post, t_atm0
        except Exception as exc:
            prepost["run_impl_tail"] = {"error": repr(exc)}

    # ---- second workload: per-column random activity, 8192 distinct columns.  The headline
    # IC above tiles 3 distinct columns (the reference's own ic::Factory::mixed), the friendliest input for class-sorted
    # kernels; this one has load imbalance in every tile, window and work queue.  Same arrays, same timing protocol.
    alt = None
    if not args.no_alt and args.ic in ("mixed", "mixed_sp"):
        try:
            atile_cols = 8192
            atile = make_tile("random", atile_cols, nlev, dtype)
            aidx = (torch.arange(ncol, device=dev) + (rank * ncol_g // world if args.scaling == "strong" else 0)) % atile_cols
            for n in FIELDS:
                t = torch.from_numpy(atile[n]).to(dev)
                fields[n].copy_(t.index_select(0, aidx))
                del t
            for n in PROGNOSTIC:
                pristine[n].copy_(fields[n])
            for _ in range(2):
                restore()
                handle.main_prepared(fstruct, infra, stream)
            barrier()
            aev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.alt_steps)]
            for s in range(args.alt_steps):
                restore()
                aev[s][0].record()
                handle.main_prepared(fstruct, infra, stream)
                aev[s][1].record()
            barrier()
            a_ms = [a.elapsed_time(b) for a, b in aev]
            a_tot = torch.tensor([float(sum(a_ms))], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(a_tot, op=dist.ReduceOp.MAX)
            a_step = float(a_tot.item()) / args.alt_steps
            aphases = handle.last_phase_ms()
            qs = 1e-14
            active = float(((atile["qc"] >= qs) | (atile["qr"] >= qs) | (atile["qi"] >= qs)).mean())
            diag = handle.get_diag(min(ncol, atile_cols))
            edges = [0, 1, 2, 4, 8, 16, 32, 64, 1 << 30]
            hist = {}
            for name, col in (("cloud", 5), ("rain", 8), ("ice", 11)):
                h, _ = np.histogram(diag[:, col], bins=edges)
                hist[name] = {"bins": "[0,1) [1,2) [2,4) [4,8) [8,16) [16,32) [32,64) [64,inf)", "columns": h.tolist(),
                              "max": int(diag[:, col].max())}
            a_ach = bpl * ncol * nlev / (float(sum(a_ms)) / args.alt_steps * 1e-3) / 1e9
            alt = {"ic": "random (G2: seeded per-column cloud layers, 8192 distinct columns)", "steps": args.alt_steps,
                   "ms_per_step": a_step, "value": ncol_g * nlev / (a_step * 1e-3), "unit": UNIT,
                   "roofline_frac": a_ach / peak, "achieved_gbs": a_ach,
                   "active_level_fraction": active, "substeps_per_column_histogram": hist,
                   "phase_ms_last_step": {k: round(v, 4) for k, v in aphases.items()},
                   "column_error_bits": handle.error_summary()}
        except Exception as exc:
            alt = {"error": repr(exc)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_oracle_rate(args.ic, nlev, dt, rb, args.cpu_seconds, predict_nc, prescribed_ccn)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f64" if rb == 8 else "f32",
                "data": "synthetic",
                "config": workload_config(args, world, ncol_g, ncol, nlev, dt), "library_flavour": args.flavour,
                "roofline": roofline, "run_impl_brackets": prepost, "conservation": budget, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
                "gpu_launches": n_launch * args.steps, "column_error_bits": err_bits,
                "step_ms": [round(x, 4) for x in step_ms],
                "phase_ms_last_step": {k: round(v, 4) for k, v in phases.items()}, "alt": alt}
        print(json.dumps(line), flush=True)
    handle.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
