B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'])"; }
$B | ph cfg3-fork
$B | ph cfg3-fork
FFEMB_SMALL_FORK=0 $B | ph cfg3-nofork
FFEMB_SMALL_FORK=0 $B | ph cfg3-nofork
