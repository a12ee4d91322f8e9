#!/bin/bash
# 2-GPU check of the bench contract (torchrun, one rank per GPU, frame-wise sharding)
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/s22_bench_n2.json 2> gpurun_out/s22_bench_n2.err
tail -c 600 gpurun_out/s22_bench_n2.err
