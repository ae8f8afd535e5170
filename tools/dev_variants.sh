one() { python bench.py --no-e2e --no-cpu --steps 4 "$@" 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['phase_ms_last_step'], d['column_error_bits'])"; }
echo base; one
for v in i5 i6 r5 r7 r4; do echo $v; P3B200_LIB=/root/repo/build/libp3b200_$v.so one; done
echo base random; one --ic random
echo i5 random; P3B200_LIB=/root/repo/build/libp3b200_i5.so one --ic random
echo i6 random; P3B200_LIB=/root/repo/build/libp3b200_i6.so one --ic random
