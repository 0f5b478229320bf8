CMD="python bench.py --rows-cap 4000000 --steps 2 --warmup 1 --no-e2e --no-cpu"
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"bwd_tiny|bwd_small|small_apply" -c 3 -o gpurun_out/r1d_small $CMD > gpurun_out/r1d_ncu.log 2>&1
tail -2 gpurun_out/r1d_ncu.log
