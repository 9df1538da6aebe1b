export ANI_B200_CHAIN=1
timeout -s KILL 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tc.py -m gpu -x -q 2>&1 | grep -v "^$" | tail -3
for b in 1 0; do
echo "== atoms 300000 chain_bwd $b"
ANI_B200_CHAIN_BWD=$b timeout -s KILL 60 python bench.py --atoms 300000 --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['stages']['mlp_ms'], d['gpu_launches_per_step'])
except Exception: print(t[-600:])"
done
echo "== trace"
ANI_B200_CHAIN_TRACE=1 timeout -s KILL 60 python bench.py --atoms 300000 --steps 1 --warmup 3 --no-cpu-baseline 2>&1 | grep -A5 "chain trace" | head -30
