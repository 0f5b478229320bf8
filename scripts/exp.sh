python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'])"; }
$B | ph cfg3-new
$B --no-overlap | ph cfg3-new-nooverlap
FFEMB_FWD_PERM=0 $B --no-overlap | ph cfg3-noperm-nooverlap
FFEMB_R1_MIX=0 $B --no-overlap | ph cfg3-nomix-nooverlap
FFEMB_FWD_PERM=0 FFEMB_R1_MIX=0 $B | ph cfg3-old
$B --workload 5 --steps 5 --warmup 2 | ph cfg5-new
FFEMB_FWD_AVG_DIV=1 $B --workload 5 --steps 5 --warmup 2 | ph cfg5-olddiv
