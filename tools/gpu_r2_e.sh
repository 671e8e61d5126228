#!/bin/bash
# 2-GPU: peer halo tests + N=2 bench (default halo fetch and the forked variant), no e2e/cpu
mkdir -p gpurun_out
python -m pytest tests/test_gpu_peer_halo.py tests/test_gpu_staging.py -m gpu -q > gpurun_out/r2_peer_v2.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r2_peer_v2.log
for f in 0; do
GRB200_HALO_FORK=$f python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$f bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_bench_n2_fork$f.json 2> gpurun_out/r2_bench_n2_fork$f.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2_bench_n2_fork$f.json'))
print('fork=$f value',round(d['value']))
for k,v in d['other_configs'].items(): print(' ',k, round(v['msps_per_gpu']), round(v['ms_per_step'],4), v['parity']['ok'])
for k,v in d['strong_scaling'].items(): print('  strong',k, round(v['msps']), round(v['ms_per_step'],4), v['parity']['ok'])
PY
done
