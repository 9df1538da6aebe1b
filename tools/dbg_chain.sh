timeout -s KILL 250 python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^$" | tail -6
for ch in 32768 65536 131072; do
echo "== chunk $ch"
ANI_B200_TC_CHUNK=$ch timeout -s KILL 80 python bench.py --atoms 300000 --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['stages']['mlp_ms'], d['gpu_launches_per_step'])
except Exception: print(t[-600:])"
done
