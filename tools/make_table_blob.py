#!/usr/bin/env python
"""Convert the reference's text ice lookup table into the binary blob the product and tests load.

Input : p3_lookup_table_1.dat-v4.1.1 (text; in the reference tree at
        components/eam/src/physics/crm/pam/external/physics/micro/p3/tables/, md5 1c959c62...).
Output: scream_b200/data/p3_lookup_table_1.v4.1.1.f64 =
        magic "P3TB" | u32 version-string length | version bytes | 12000 f64 ice | 60000 f64 coll_raw
        (little-endian).  Values are the exact doubles `in >> double` yields; the collection
        values are stored RAW (log10 is applied at table-init time like p3_table_ice_impl.hpp:63).
Parse rules follow p3_table_ice_impl.hpp:33-71.  Run once in the build container; the GPU box has
no /root/reference, so the blob is what travels.
"""
import struct
import sys
import numpy as np

SRC = ("/root/reference/components/eam/src/physics/crm/pam/external/physics/micro/p3/tables/"
       "p3_lookup_table_1.dat-v4.1.1")


def parse(path):
    toks = open(path).read().split()
    assert toks[0] == "VERSION", toks[:2]
    version = toks[1]
    pos = 2
    ice = np.zeros((5, 4, 50, 12))
    coll = np.zeros((5, 4, 50, 30, 2))
    for jj in range(5):
        for ii in range(4):
            for i in range(50):
                row = toks[pos:pos + 17]; pos += 17
                vals = [float(x) for x in row[2:]]
                ice[jj, ii, i, :] = [v for j, v in enumerate(vals) if j > 1 and j != 10]
            for i in range(50):
                for j in range(30):
                    row = toks[pos:pos + 8]; pos += 8
                    vals = [float(x) for x in row[2:]]
                    coll[jj, ii, i, j, :] = vals[3:5]
    assert pos == len(toks), (pos, len(toks))
    return version, ice, coll


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else SRC
    out = sys.argv[2] if len(sys.argv) > 2 else "scream_b200/data/p3_lookup_table_1.v4.1.1.f64"
    version, ice, coll = parse(src)
    assert (coll > 0).all()
    with open(out, "wb") as f:
        v = version.encode()
        f.write(b"P3TB" + struct.pack("<I", len(v)) + v)
        f.write(ice.astype("<f8").tobytes())
        f.write(coll.astype("<f8").tobytes())
    print(out, version, ice.shape, coll.shape)


if __name__ == "__main__":
    main()
