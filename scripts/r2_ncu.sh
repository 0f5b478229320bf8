# round 2 ncu evidence (one GPU, run under gpurun).  usage: r2_ncu.sh TAG
#  1. plain runs first (exit 0 required)
#  2. cfg3 AT THE BENCHED SIZE (96 GB of tables): every kernel of the timed region with its duration and
#     DRAM bytes — application replay (the program is re-run per pass), so no save/restore of the tables
#  3. --set full of our kernels at rows capped to 4M (kernel replay needs to save/restore what a kernel writes)
#  4. TMA A/B: forward with LDG rows (product) vs cp.async.bulk rows (FFEMB_FWD_BULK=S), timing without ncu,
#     then the counters of both under ncu
TAG=${1:-r2n}
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"
FULL="python bench.py --steps 2 --warmup 1 --pool 4 --no-e2e --no-cpu"
$FULL > gpurun_out/${TAG}_plain_full.log 2>&1 || { echo "plain full-size run failed"; tail -5 gpurun_out/${TAG}_plain_full.log; exit 1; }
# parity of the bulk-copy forward before it is measured
FFEMB_FWD_BULK=8 python -m pytest tests/test_gpu_parity.py -q -x -k "forward_bit_exact or forward_golden" > gpurun_out/${TAG}_bulk_parity.log 2>&1; echo "bulk forward parity rc=$?"; tail -2 gpurun_out/${TAG}_bulk_parity.log
for S in 0 4 8 16; do
  FFEMB_FWD_BULK=$S python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu --no-overlap > gpurun_out/${TAG}_fwd_bulk$S.json 2> gpurun_out/${TAG}_fwd_bulk$S.err
  python -c "
import json
d=json.loads(open('gpurun_out/${TAG}_fwd_bulk$S.json').read().strip().splitlines()[-1])
print('FFEMB_FWD_BULK=$S cfg3 step ms', round(d['ms_per_step'],4), 'forward ms', d['roofline']['phases_ms_per_step'].get('forward'))"
done
ncu --replay-mode application --profile-from-start off --metrics $M --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_full_size_launches.csv $FULL > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu full-size rc=$?"
CAP="python bench.py --rows-cap 4000000 --steps 2 --warmup 1 --no-e2e --no-cpu"
$CAP > gpurun_out/${TAG}_plain_cap.log 2>&1 && ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"fwd_kernel|reduce1_|combine_kernel|keygen_kernel|bwd_small|bwd_tiny|small_apply" -c 9 -o gpurun_out/${TAG}_full $CAP > gpurun_out/${TAG}_ncu2.log 2>&1; echo "ncu set-full rc=$?"
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
FFEMB_FWD_BULK=8 $FULL > gpurun_out/${TAG}_plain_bulk.log 2>&1 && FFEMB_FWD_BULK=8 ncu --replay-mode application --profile-from-start off --metrics $M --clock-control none -k regex:"fwd_bulk" -c 2 --csv --log-file gpurun_out/${TAG}_fwd_bulk8_ncu.csv $FULL > gpurun_out/${TAG}_ncu3.log 2>&1; echo "ncu bulk rc=$?"
for f in gpurun_out/${TAG}_*.ncu-rep; do [ -f "$f" ] && [ $(stat -c %s "$f") -gt 30000000 ] && rm -f "$f"; done
wc -l gpurun_out/${TAG}_full_size_launches.csv gpurun_out/${TAG}_fwd_bulk8_ncu.csv; du -sh gpurun_out
