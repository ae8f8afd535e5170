"""Debug helper (GPU box): random-column parity dump, GPU vs oracle vs the 1-lane host build."""
import sys
import numpy as np
sys.path.insert(0, ".")
from scream_b200 import ic_cases, lib
from scream_b200.fields import OUTPUTS
from tests import oracle_lib, parity, hostsim_lib

nlev = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dt = float(sys.argv[2]) if len(sys.argv) > 2 else 300.0
h = lib.P3Handle(0, 8)
f = ic_cases.make_random(4096, nlev)
ref, got, d_ref, d_got = parity.run_pair(oracle_lib, h, f, dt, True, True)
sim = parity.clone(f); parity.zero_outputs(sim)
d_sim = hostsim_lib.p3_main(sim, dt, True, True)
print("oracle vs hostsim: cols>1e-12", parity.column_mismatch(ref, sim, 1e-12).sum(), "diag", (d_ref != d_sim).any(axis=1).sum())
bad = parity.column_mismatch(ref, got, 1e-12)
dbad = (d_ref != d_got).any(axis=1)
print("oracle vs gpu: cols>1e-12", bad.sum(), "diag", dbad.sum())
for c in np.argwhere(dbad)[:8, 0]:
    print("col", c, "\n  ref", d_ref[c, :13].tolist(), "\n  gpu", d_got[c, :13].tolist())
cols = np.argwhere(bad & ~dbad)[:8, 0]
for c in cols:
    print("col", c, "diag", d_ref[c, :13].tolist())
    for n in OUTPUTS:
        x = np.atleast_1d(ref[n][c]); y = np.atleast_1d(got[n][c])
        den = np.abs(ref[n]).max() or 1.0
        d = np.abs(x - y) / den
        if d.max() > 1e-12:
            k = int(d.argmax())
            print("   ", n, "k", k, "ref", x[k], "gpu", y[k], "err", d.max(), "levels", np.argwhere(d > 1e-12)[:10, 0].tolist())
# per-field worst
e = parity.field_errors(ref, got)
print({k: float("%.3g" % v) for k, v in e.items() if v > 1e-13})
for n in ("liq_ice_exchange", "vap_liq_exchange", "vap_ice_exchange", "qv2qi_depos_tend", "nc", "qc", "th"):
    x = ref[n]; y = got[n]
    den = np.abs(x).max() or 1.0
    d = np.abs(x - y) / den
    idx = np.argwhere(d > 1e-12)
    print(n, "n_bad", len(idx), "cols", sorted(set(idx[:, 0].tolist()))[:10])
    for c, k in idx[:6]:
        print("    ", (int(c), int(k)), x[c, k], y[c, k], "T", f["th"][c, k] / f["inv_exner"][c, k], "qc qr qi", f["qc"][c, k], f["qr"][c, k], f["qi"][c, k])
