for ov in 0 1 0 1; do
ANI_BENCH_HALO_OVERLAP=$ov ANI_BENCH_NO_CPU=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 60 --warmup 5 --no-md > gpurun_out/halo8_$ov.json 2> gpurun_out/halo8_$ov.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/halo8_$ov.json").read().strip().splitlines()[-1]); print("N=8 overlap=$ov", d["value"], d["ms_per_step"], d.get("multi_gpu_force_check"))
except Exception as e: print("ERR", e, open("gpurun_out/halo8_$ov.err").read()[-1500:])
PY
done
