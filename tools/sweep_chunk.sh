for ch in 2048 4096 8192 16384 32768; do
  ANI_B200_TC_CHUNK=$ch timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('chunk',$ch,'ms',round(d['ms_per_step'],3),'mlp',round(d['stages']['mlp_ms'],3),'launches',d['gpu_launches_per_step'])"
done
for ch in 4096 8192; do
  ANI_B200_PIPE=1 ANI_B200_TC_CHUNK=$ch timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('PIPE chunk',$ch,'ms',round(d['ms_per_step'],3),'mlp',round(d['stages']['mlp_ms'],3),'launches',d['gpu_launches_per_step'])"
done
