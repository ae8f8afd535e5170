"""Experiment: does p3_main gain from running column chunks side by side (one handle + stream per lane), so that the
HBM-bound passes of one chunk (classification, phase C) run under the issue-bound sedimentation of another?
Usage: python tools/exp_chunk_overlap.py [--ncol N] [--steps S]      (GPU box only; prints one JSON line per layout)
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from scream_b200.fields import FIELDS, OUTPUTS, PROGNOSTIC  # noqa: E402


def main():
    import torch
    import bench
    from scream_b200 import lib as p3
    ap = argparse.ArgumentParser()
    ap.add_argument("--ncol", type=int, default=1572864)
    ap.add_argument("--nlev", type=int, default=128)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--ic", default="mixed")
    ap.add_argument("--layouts", default="1x1,2x2,2x4,2x8,3x6,4x8,4x16")
    ap.add_argument("--priority", type=int, default=0)
    ap.add_argument("--stagger", type=int, default=0)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    ncol, nlev, dt = args.ncol, args.nlev, 300.0
    tile_cols = 3 if args.ic.startswith("mixed") else 8192
    tile = bench.make_tile(args.ic, tile_cols, nlev, np.float64)
    idx = torch.arange(ncol, device=dev) % tile_cols
    fields = {n: torch.from_numpy(tile[n]).to(dev).index_select(0, idx).contiguous() for n in FIELDS}
    pristine = {n: fields[n].clone() for n in PROGNOSTIC}
    zero_names = [n for n in OUTPUTS if n not in PROGNOSTIC]

    def restore():
        for n in PROGNOSTIC:
            fields[n].copy_(pristine[n])
        for n in zero_names:
            fields[n].zero_()

    ref = None
    for layout in args.layouts.split(","):
        K, C = (int(x) for x in layout.split("x"))
        handles = [p3.P3Handle(0, 8) for _ in range(K)]
        streams = [torch.cuda.Stream(priority=args.priority) for _ in range(K)]
        chunks = []
        if not args.stagger:
            bounds = [ncol * c // C for c in range(C + 1)]
            plan = [(bounds[c], bounds[c + 1], c % K) for c in range(C)]
        else:
            # lane l owns columns [l/K, (l+1)/K) of the grid, in C/K chunks whose boundaries are shifted by l/K of a chunk
            plan = []
            per = C // K
            for c in range(per + 1):
                for l in range(K):
                    base, span = ncol * l // K, ncol // K
                    edges = [0] + [min(span, int(span * (i - l / K) / per)) for i in range(1, per + 1)] + [span]
                    edges = sorted(set(max(0, e) for e in edges))
                    if c < len(edges) - 1:
                        plan.append((base + edges[c], base + edges[c + 1], l))
        for lo, hi, l in plan:
            sub = {n: fields[n][lo:hi] for n in FIELDS}
            h = handles[l]
            f, nc_, nk_ = h.fields_struct(sub, host=False)
            chunks.append((h, f, h._infra(nc_, nk_, dt, True, True, True), streams[l], sub))
        main_stream = torch.cuda.current_stream()
        times = []
        for s in range(args.steps + 2):
            restore()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main_stream)
            for st in streams:
                st.wait_event(e0)
            for h, f, infra, st, _ in chunks:
                h.main_prepared(f, infra, st.cuda_stream)
            for st in streams:
                done = torch.cuda.Event()
                done.record(st)
                main_stream.wait_event(done)
            e1.record(main_stream)
            torch.cuda.synchronize()
            if s >= 2:
                times.append(e0.elapsed_time(e1))
        out = {n: fields[n].clone() for n in ("qc", "qr", "qi", "qv", "th", "precip_liq_surf", "precip_ice_surf")}
        same = None
        if ref is None:
            ref = out
        else:
            same = all(torch.equal(ref[n].view(torch.int64), out[n].view(torch.int64)) for n in ref)
        print(json.dumps({"lanes": K, "chunks": C, "priority": args.priority, "stagger": args.stagger, "plan": [(hi - lo, l) for lo, hi, l in plan][:8], "ms": [round(t, 3) for t in times],
                          "ms_min": round(min(times), 3), "bit_identical_to_first_layout": same}), flush=True)
        for h in handles:
            h.close()


if __name__ == "__main__":
    main()
