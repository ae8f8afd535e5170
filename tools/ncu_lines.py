#!/usr/bin/env python
"""Attribute an `ncu --page source --csv` SASS profile to source lines of the product code.

usage: tools/ncu_lines.py <report.ncu-rep> <cubin> <kernel-mangled-substring> [--top N] [--launch <demangled-substring>]

Joins ncu's per-SASS-instruction counters (instructions executed, stall samples) with the line info
`nvdisasm -c -gi` prints for the same cubin (built with -lineinfo), by instruction offset.  Every
instruction is attributed (a) to the outermost frame in p3_column.cuh (which phase of the column
driver), and (b) to the innermost frame inside our sources (which physics statement).
"""
import collections
import csv
import io
import re
import subprocess
import sys


def disasm(cubin, kernel):
    txt = subprocess.run(["nvdisasm", "-c", "-gi", cubin], capture_output=True, text=True).stdout
    out = {}            # offset -> list of (file, line) innermost first
    cur = []
    pending = []
    infunc = False
    sub = None
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            infunc = kernel in ln
            continue
        if not infunc:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            pending.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m = re.match(r"\s*\$?([\w$]+):\s*$", ln)
        if m and "$" in ln:
            sub = m.group(1).split("$")[-1]
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*);", ln)
        if m:
            if pending:
                cur = pending
                pending = []
            off = int(m.group(1), 16)
            out[off] = (list(cur), m.group(2).strip(), sub)
    return out


def main():
    rep, cubin, kernel = sys.argv[1:4]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    if rep.endswith(".csv") or rep.endswith(".csv.gz"):   # `ncu -i <rep> --page source --csv`, exported on the GPU box (tools/ncu_export.sh)
        import gzip
        raw = (gzip.open(rep, "rt") if rep.endswith(".gz") else open(rep)).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    lines = raw.splitlines()
    # one section per profiled launch: "Kernel Name",<demangled> then the "Address",... table
    want = sys.argv[sys.argv.index("--launch") + 1] if "--launch" in sys.argv else None
    heads = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
    if heads and want is not None:
        h = next(i for i in heads if want in lines[i])
        end = next((j for j in heads if j > h), len(lines))
        lines = lines[h:end]
        print("launch:", lines[0][:120])
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
    end = next((j for j in range(start + 1, len(lines)) if lines[j].startswith('"Kernel Name"')), len(lines))
    rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:end]))))
    base = int(rows[0]["Address"], 16)
    dis = disasm(cubin, kernel)
    by_phase = collections.Counter()
    by_phase_s = collections.Counter()
    by_inner = collections.Counter()
    by_inner_s = collections.Counter()
    by_sub = collections.Counter()
    tot_i = tot_s = 0
    stall_cols = [c for c in rows[0].keys() if c.startswith("stall_") and "Not Issued" not in c]
    stall_by_phase = collections.defaultdict(collections.Counter)
    for r in rows:
        off = int(r["Address"], 16) - base
        n = int(r["Instructions Executed"] or 0)
        s = int(r["# Samples"] or 0)
        tot_i += n
        tot_s += s
        frames, text, sub = dis.get(off, ([], "", None))
        phase = next(((f, l) for f, l in reversed(frames) if f == "p3_column.cuh"), None)
        inner = next(((f, l) for f, l in frames if f.startswith("p3_")), None)
        if sub and not frames:
            phase = inner = ("<" + sub + ">", 0)
        by_sub[sub or "inline"] += n
        by_phase[phase] += n
        by_phase_s[phase] += s
        by_inner[inner] += n
        by_inner_s[inner] += s
        for c in stall_cols:
            v = int(r[c] or 0)
            if v:
                stall_by_phase[phase][c] += v
    print(f"total warp instructions {tot_i:,}  samples {tot_s:,}")
    if "--calls" in sys.argv:
        # every CALL that was executed: a call site of an out-of-line routine (fp64 division / reciprocal slow paths,
        # pow) counts the warps that left the inline fast path there
        print("\n== executed CALL sites ==")
        sites = []
        for r in rows:
            off = int(r["Address"], 16) - base
            frames, text, sub = dis.get(off, ([], "", None))
            n = int(r["Instructions Executed"] or 0)
            if "CALL" in text and n:
                sites.append((n, off, text, frames))
        for n, off, text, frames in sorted(sites, reverse=True)[:top]:
            fr = " <- ".join(f"{f}:{l}" for f, l in frames[:4])
            print(f"  {n:12,d}  {text[:60]:60s} {fr}")
    print("\n== by device subroutine (called, not inlined) ==")
    for k, v in by_sub.most_common(12):
        print(f"  {str(k):40s} {100*v/tot_i:6.2f}% inst")
    print("\n== by column-driver line (outermost p3_column.cuh frame) ==")
    for k, v in by_phase_s.most_common(top):
        st = ", ".join(f"{c[6:]}={100*x/max(v,1):.0f}%" for c, x in stall_by_phase[k].most_common(3))
        print(f"  {str(k):32s} {100*v/tot_s:6.2f}% samples  {100*by_phase[k]/tot_i:6.2f}% inst   {st}")
    print("\n== by innermost product-source line ==")
    for k, v in by_inner_s.most_common(top):
        print(f"  {str(k):32s} {100*v/tot_s:6.2f}% samples  {100*by_inner[k]/tot_i:6.2f}% inst")


if __name__ == "__main__":
    main()
