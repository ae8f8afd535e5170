timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-e2e --no-cpu --steps 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['run_impl_brackets'])"
