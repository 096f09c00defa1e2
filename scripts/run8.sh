mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29701 scripts/bench_allreduce.py > gpurun_out/allreduce8b.txt 2>&1; grep case gpurun_out/allreduce8b.txt | cut -c1-200
timeout 200 $TR --master-port 29703 bench.py --gpus 8 --steps 30 --warmup 5 2>&1 | tail -1 | tee gpurun_out/bench8b.txt | cut -c1-160
timeout 200 $TR --master-port 29704 bench.py --gpus 8 --steps 30 --warmup 5 --comm nccl 2>&1 | tail -1 | tee -a gpurun_out/bench8b.txt | cut -c1-160
timeout 300 $TR --master-port 29702 scripts/bench_fed.py --mode all --rounds 2 > gpurun_out/fed8b.txt 2>&1; grep bench gpurun_out/fed8b.txt | cut -c1-330
