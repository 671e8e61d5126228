#!/bin/bash
# two GPUs: the peer-halo tests (skipped on one GPU) and the bench line at N = 2 as the driver launches it
mkdir -p gpurun_out
nvidia-smi -L | head -4
timeout 900 python -m pytest tests -m gpu -x -q -k "peer or halo or shard or multi" > gpurun_out/r2_gpu_tests_n2_v2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_gpu_tests_n2_v2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2_v2.json 2> gpurun_out/r2_bench_n2_v2.err; echo "bench n=2 rc=$?"
tail -c 400 gpurun_out/r2_bench_n2_v2.err
