python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu --no-overlap"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'])"; }
$B | ph cfg3
$B | ph cfg3
$B --workload 5 --steps 5 --warmup 2 | ph cfg5
$B --workload 1 | ph cfg1
