#!/bin/bash
# Run on the GPU box: one `ncu --set full` capture, exported as the two CSV pages the tools here read
# (tools/ncu_summary.py reads raw.csv, tools/ncu_lines.py reads source.csv); the .ncu-rep itself (tens of MB with
# --import-source) is deleted so that gpurun_out stays under its size limit.
# usage: tools/ncu_export.sh <out-prefix> <ncu kernel filter args...> -- <command...>
set -u
out=$1; shift
filt=()
while [ "$1" != "--" ]; do filt+=("$1"); shift; done
shift
ncu --set full --clock-control none --import-source on "${filt[@]}" -f -o /tmp/ncu_export "$@" > "${out}_ncu.log" 2>&1
rc=$?
ncu -i /tmp/ncu_export.ncu-rep --page raw --csv > "${out}_raw.csv" 2>/dev/null
ncu -i /tmp/ncu_export.ncu-rep --page source --csv > "${out}_source.csv" 2>/dev/null
gzip -f "${out}_source.csv"
rm -f /tmp/ncu_export.ncu-rep
exit $rc
