python -m pytest tests -m gpu -x -q 2>&1 | tail -2
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu --no-overlap"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'])"; }
$B | ph cfg3
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -k regex:"bwd_tiny|bwd_small|small_apply" -c 3 --csv python bench.py --rows-cap 4000000 --steps 2 --warmup 1 --no-e2e --no-cpu 2>/dev/null | grep -E "bwd_|small_apply" | awk -F'","' '{print $5, $(NF-1), $NF}'
