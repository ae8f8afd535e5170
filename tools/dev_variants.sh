one() { python bench.py --no-e2e --no-cpu --steps 4 "$@" 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['phase_ms_last_step'], d['column_error_bits'])"; }
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo base-mixed; one
echo base-random; one --ic random
