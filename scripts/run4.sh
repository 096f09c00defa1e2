mkdir -p gpurun_out
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
TR2="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR4 --master-port 29711 bench.py --gpus 4 --steps 30 --warmup 5 2>&1 | tail -1 | tee gpurun_out/scale4.txt | cut -c1-170
timeout 200 $TR2 --master-port 29712 bench.py --gpus 2 --steps 30 --warmup 5 2>&1 | tail -1 | tee -a gpurun_out/scale4.txt | cut -c1-170
timeout 200 $TR4 --master-port 29713 baseline/ddp_baseline.py --gpus 4 --steps 20 --warmup 5 2>&1 | tail -1 | tee -a gpurun_out/scale4.txt | cut -c1-250
timeout 200 $TR2 --master-port 29714 baseline/ddp_baseline.py --gpus 2 --steps 20 --warmup 5 2>&1 | tail -1 | tee -a gpurun_out/scale4.txt | cut -c1-250
