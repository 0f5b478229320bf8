TAG=$1; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest.log
run() { name=$1; shift
  env "$@" python bench.py --no-cpu --no-e2e $ARGS > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err; rc=$?
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_$name.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('$name', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'launches/step', d['gpu_launches']/d['steps'], r.get('phases_ms_per_step'))
except Exception as e:
    print('$name failed rc=$rc', e); print(open('gpurun_out/${TAG}_$name.err').read()[-1500:])
PY
}
ARGS="--workload 2 --steps 20 --warmup 5" run cfg2 A=1
ARGS="--workload 2 --steps 20 --warmup 5" run cfg2_onecta FFEMB_OWN_SORT=2
ARGS="--workload 1 --steps 20 --warmup 5" run cfg1 A=1
