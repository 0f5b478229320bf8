# round 2, first single-GPU call: parity suite (incl. the full-size tests), smoke, bench cfg3/1/2/5, reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2z_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/r2z_cfg3.json 2> gpurun_out/r2z_cfg3.err; echo "cfg3 rc=$?"
for w in 1 2 5; do python bench.py --workload $w > gpurun_out/r2z_cfg$w.json 2> gpurun_out/r2z_cfg$w.err; echo "cfg$w rc=$?"; done
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2z_ref.json 2> gpurun_out/r2z_ref.err; echo "ref rc=$?"
for f in cfg3 cfg1 cfg2 cfg5 ref; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2z_$f.json').read().strip().splitlines()[-1])
    r=d.get('roofline',{})
    print('$f', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), r.get('kernel'), r.get('frac'), r.get('frac_8d'), r.get('frac_compulsory'), r.get('phases_ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_step'), 'cpu', (d.get('cpu_baseline') or {}).get('value'), (d.get('cpu_baseline') or {}).get('by_threads'), 'refgpu', (d.get('reference_on_b200') or {}).get('ms_per_step'))
except Exception as e:
    print('$f', 'failed', e); print(open('gpurun_out/r2z_$f.err').read()[-1500:])
PY
done
cat gpurun_out/r2_parity_fullsize.json 2>/dev/null | head -80
