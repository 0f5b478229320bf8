mkdir -p gpurun_out
set -x
CMD="python bench.py --rows-cap 4000000 --steps 2 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fwd_pool_kernel|reduce_level_kernel|keygen" -c 5 -o gpurun_out/prof_r1a $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/plain.log; tail -5 gpurun_out/ncu2.log
