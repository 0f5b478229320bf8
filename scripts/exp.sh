python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "host_step" 2>&1 | tail -3
for w in 1 2 3; do python bench.py --workload $w --steps 20 --warmup 5 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cfg$w', round(d['ms_per_step'],3), d['e2e'])"; done
