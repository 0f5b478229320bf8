TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
mkdir -p gpurun_out
timeout 300 $TR tests/multigpu_check.py > gpurun_out/r2w_mgcheck_n2.log 2>&1; echo "multigpu_check rc=$?"; grep -i "assert\|ok:\|Error" gpurun_out/r2w_mgcheck_n2.log | head -5
