one() { python bench.py --no-e2e --no-cpu --steps 4 "$@" 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['phase_ms_last_step']['phase_a'], d['phase_ms_last_step']['phase_c'], d['column_error_bits'])"; }
echo base; one
for v in c256m3 c256m2 c128m6 c128m8 c256m5; do echo $v; P3B200_LIB=/root/repo/build/libp3b200_$v.so one; done
