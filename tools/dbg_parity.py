"""Debug helper (GPU box): per-flag parity dump of the mixed IC through the C ABI vs the oracle."""
import sys
import numpy as np
sys.path.insert(0, ".")
from scream_b200 import ic_cases, lib
from tests import oracle_lib, parity

rb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
h = lib.P3Handle(0, rb)
dtype = np.float64 if rb == 8 else np.float32
for flags in [(False, False), (False, True), (True, False), (True, True), (False, False)]:
    f = ic_cases.make_mixed(3, 72, dtype)
    ref, got, d_ref, d_got = parity.run_pair(oracle_lib, h, f, 300.0, *flags)
    e = parity.field_errors(ref, got)
    print(flags, "diag equal", np.array_equal(d_ref, d_got), "worst", max(e.values()))
    if not np.array_equal(d_ref, d_got):
        print(d_ref[:, :13]); print(d_got[:, :13])
    for k, v in sorted(e.items(), key=lambda x: -x[1])[:6]:
        x = ref[k]; y = got[k]; diff = np.abs(x - y)
        bad = np.argwhere(diff > parity.TOL[rb] * np.abs(x).max())
        print("   ", k, v, bad[:8].tolist())
        for idx in bad[:3]:
            print("       ", tuple(idx), x[tuple(idx)], y[tuple(idx)])
