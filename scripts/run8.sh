mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gpu_multi.py -q -x 2>&1 | tail -3
timeout 200 $TR --master-port 29701 scripts/bench_allreduce.py > gpurun_out/allreduce8.txt 2>&1; tail -12 gpurun_out/allreduce8.txt
timeout 300 $TR --master-port 29702 scripts/bench_fed.py --mode all --rounds 3 > gpurun_out/fed8.txt 2>&1; grep bench gpurun_out/fed8.txt
for args in "" "--no-overlap" "--comm nccl"; do timeout 200 $TR --master-port 29703 bench.py --gpus 8 --steps 30 --warmup 5 $args 2>&1 | tail -1 | tee -a gpurun_out/bench8.txt | cut -c1-160; done
for m in vgg16 mobilenet_v2; do for args in "" "--comm nccl"; do timeout 200 $TR --master-port 29704 bench.py --gpus 8 --model $m --steps 30 --warmup 5 $args 2>&1 | tail -1 | tee -a gpurun_out/bench8.txt | cut -c1-160; done; done
