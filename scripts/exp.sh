#!/bin/bash
# scratch: launch lists of the small configs (which kernels make up the graph-replayed step)
mkdir -p gpurun_out
for w in 1 2; do
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/l_cfg${w}.csv python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu --graph off > gpurun_out/l_cfg${w}.log 2>&1
done
python - <<'P'
import csv,re
for w in (1,2):
    rows=list(csv.reader(open(f'gpurun_out/l_cfg{w}.csv')))
    hdr=None;tot=0
    for r in rows:
        if 'Kernel Name' in r: hdr=r; continue
        if hdr and len(r)==len(hdr):
            d=dict(zip(hdr,r))
            if d.get('Metric Name')!='gpu__time_duration.sum': continue
            v=float(d['Metric Value'].replace(',',''))/1e3
            if int(d['ID'])<14: print(w,d['ID'],re.sub(r'\(.*','',d['Kernel Name'])[:70],d['Grid Size'],d['Block Size'],round(v,1))
            tot+=v
    print('total us for 2 steps',round(tot,1))
P
