#!/bin/bash
# final ncu captures: conv kernels at block-1 and block-3 shapes, the three comm kernels (8 ranks emulated on one GPU)
set -o pipefail
for h in 13 3; do
python scripts/ncu_small.py $h > gpurun_out/plain_small_$h.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_ -s 6 -c 6 -f -o gpurun_out/final_conv_h$h python scripts/ncu_small.py $h > gpurun_out/ncu_final_h$h.log 2>&1
echo "ncu conv h=$h rc=$?"
done
python scripts/profile_comm.py > gpurun_out/plain_comm.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'allreduce_optim|fedavg_reduce|secagg' -c 12 -f -o gpurun_out/final_comm python scripts/profile_comm.py > gpurun_out/ncu_final_comm.log 2>&1
echo "ncu comm rc=$?"; tail -2 gpurun_out/plain_comm.log
