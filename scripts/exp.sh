python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "forward" 2>&1 | tail -2
B="python bench.py --steps 5 --warmup 2 --no-e2e --no-cpu --no-overlap"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'])"; }
$B --workload 5 | ph cfg5
