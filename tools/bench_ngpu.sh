N=$1
export FB_SHARD_TIMING=1
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
echo rc=$?
grep -E "fb shard" gpurun_out/r2_bench_${N}gpu.err | sort | uniq -c | sort -rn | head -12
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_${N}gpu.json') if l.startswith('{')][-1])
for k in ('n_gpus','value','ms_per_step','e2e','c3_sweep','c4_trailblock','single_process_2gpu_bit_identical'):
    print(k, d.get(k))
c=d.get('c5_chromosome'); 
if c: c['admm_iters_per_solve']=c['admm_iters_per_solve'][:6]; print(c)
PY
