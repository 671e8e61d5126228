#!/bin/bash
# 2-GPU call: peer-halo test, staging tests (device-scope test needs 2 GPUs), bench at N=2
mkdir -p gpurun_out
python -m pytest tests/test_gpu_peer_halo.py tests/test_gpu_staging.py -m gpu -q > gpurun_out/r2_peer_v1.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_peer_v1.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_bench_n2_v1.json 2> gpurun_out/r2_bench_n2_v1.err; echo "bench rc=$?"
tail -5 gpurun_out/r2_bench_n2_v1.err
