nvidia-smi -L
python -m pytest tests/test_gpu_dist.py -q -rs 2>&1 | tail -5
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 6 --warmup 2 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err; echo rc=$?
tail -c 1500 gpurun_out/r02_bench_2gpu.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_2gpu.json').read().strip().splitlines()[-1])
for k in ("value","n_gpus","ms_per_step","e2e","d16","c5","c4","sharded","clocks"):
    print(k, json.dumps(d.get(k))[:700])
P
