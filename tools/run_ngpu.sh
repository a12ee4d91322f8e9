#!/bin/bash
# N-GPU check of the bench contract (torchrun, one rank per GPU, frame-wise sharding): tools/run_ngpu.sh N
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -c 400 gpurun_out/bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/ref_n$N.json 2> gpurun_out/ref_n$N.err
tail -c 300 gpurun_out/ref_n$N.err
