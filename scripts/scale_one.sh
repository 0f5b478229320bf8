# one point of the scaling curve: bench.py at N = $1 GPUs (fused exchange, balanced placement), run on a
# box with exactly N GPUs (gpurun --gpus N) so that the charge is N x the box time of this point only.
# N = 8 also runs the multi-GPU parity check and BASELINE config 4 (row-wise, 51 GB of tables per GPU).
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
$TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > gpurun_out/r1f_n$N.json 2> gpurun_out/r1f_n$N.err; echo "bench N=$N rc=$?"
if [ "$N" = "8" ]; then
  FFEMB_FWD_PERM=0 $TR bench.py --gpus 8 --steps 30 --warmup 5 --no-cpu --no-e2e > gpurun_out/r1f_n8_noperm.json 2> gpurun_out/r1f_n8_noperm.err; echo "noperm rc=$?"
  python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -2
  $TR bench.py --gpus 8 --workload 4 --steps 20 --warmup 5 > gpurun_out/r1f_cfg4_n8.json 2> gpurun_out/r1f_cfg4_n8.err; echo "cfg4 rc=$?"
fi
for f in n$N n8_noperm cfg4_n8; do [ ! -f gpurun_out/r1f_$f.json ] || python -c "
import json
d=json.loads(open('gpurun_out/r1f_$f.json').read().strip().splitlines()[-1])
print('$f', 'ms', round(d['ms_per_step'],3), 'Glookups/s', round(d['value']/1e9,2), d['roofline'].get('phases_ms_per_step'), 'e2e', d.get('e2e',{}).get('ms_per_step'), d['config'].get('table_owner'))
"; done
