timeout 100 python tools/fir_tc_check.py 2>&1 | grep "tc=1" | head -3
for i in 1 2; do timeout 100 python bench.py --workload fir --steps 20 --warmup 5 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['frac'],3))"; done
