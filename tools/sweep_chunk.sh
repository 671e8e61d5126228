#!/bin/bash
# e2e (host-buffer C-ABI) throughput of the channelizer vs the staging granularity GRB200_CHUNK_MB
for mb in 16 32 64 128 256 512; do
  r=$(GRB200_CHUNK_MB=$mb timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-others 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['e2e']['value']), round(d['e2e']['pcie_gbs_each_way'],1))")
  echo "chunk_mb $mb e2e_msps/pcie $r"
done
