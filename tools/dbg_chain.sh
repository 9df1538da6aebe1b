timeout -s KILL 200 python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^$" | tail -4
for n in 100000 300000; do for b in 1 0; do
echo "== atoms $n chain_bwd $b"
ANI_B200_CHAIN_BWD=$b timeout -s KILL 60 python bench.py --atoms $n --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['stages'], d['gpu_launches_per_step'])
except Exception: print(t[-600:])"
done; done
