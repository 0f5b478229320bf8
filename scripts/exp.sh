FFEMB_R1_RING=8 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
FFEMB_R1_RING=4 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "backward" 2>&1 | tail -2
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu --no-overlap"
ph() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['roofline']['phases_ms_per_step'])"; }
$B | ph regs
FFEMB_R1_RING=4 $B | ph ring4
FFEMB_R1_RING=8 $B | ph ring8
FFEMB_R1_RING=8 $B --workload 5 | ph ring8_cfg5
$B --workload 5 | ph regs_cfg5
