# ncu evidence for one round (run under gpurun, one GPU).  $1 = tag (e.g. r1b).
# 1) plain run must exit 0;  2) launch list (gpu__time_duration) of the same command, timed region
# only (bench.py brackets it with cudaProfilerStart/Stop);  3) one --set full capture of our kernels.
# Rows are capped at 4M for the replays (ncu saves/restores written memory per pass; the 96 GB
# working set would take minutes per kernel).
TAG=${1:-r1}
mkdir -p gpurun_out
CMD="python bench.py --rows-cap 4000000 --steps 2 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"fwd_kernel|reduce1_|keygen_kernel|reduce_level_kernel|reduce_tail_kernel|bwd_small|bwd_tiny|small_apply" -c 11 -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
tail -1 gpurun_out/${TAG}_plain.log | cut -c1-300; tail -3 gpurun_out/${TAG}_ncu2.log; wc -l gpurun_out/${TAG}_launches.csv
# cfg5 (Zipf, AVG, bag 64) forward only, full size: which unit bounds it under hot-row skew
if [ -n "$CFG5" ]; then
  C5="python bench.py --workload 5 --steps 1 --warmup 1 --no-e2e --no-cpu --no-overlap"
  ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"fwd_kernel" -c 1 -o gpurun_out/${TAG}_cfg5fwd $C5 > gpurun_out/${TAG}_ncu3.log 2>&1
  ncu -i gpurun_out/${TAG}_cfg5fwd.ncu-rep --page raw --csv > gpurun_out/${TAG}_cfg5fwd_raw.csv 2>/dev/null
  tail -2 gpurun_out/${TAG}_ncu3.log
fi
# gpurun brings back at most 64 MiB: the raw CSV pages are what gets read; keep a report only if small
for f in gpurun_out/${TAG}_*.ncu-rep; do [ -f "$f" ] && [ $(stat -c %s "$f") -gt 20000000 ] && rm -f "$f"; done
du -sh gpurun_out
