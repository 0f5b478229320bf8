python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu --no-overlap"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'], d.get('e2e'))"; }
python bench.py --steps 20 --warmup 5 --no-cpu | ph cfg3-e2e
for kb in 1024 4096 32768; do FFEMB_L1_KB=$kb $B | ph cfg3-l1kb-$kb; done
$B --workload 5 --steps 5 --warmup 2 | ph cfg5
FFEMB_L1_KB=100000000 $B --workload 5 --steps 5 --warmup 2 | ph cfg5-l1-all
