nproc; free -g | head -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus 8 --steps 10 --warmup 3 --quick > gpurun_out/r02_bench_8gpu_quick.json 2> gpurun_out/r02_bench_8gpu_quick.err; echo rc=$?
tail -c 600 gpurun_out/r02_bench_8gpu_quick.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_8gpu_quick.json').read().strip().splitlines()[-1])
print(d["value"], d["n_gpus"], d["ms_per_step"], d["e2e"]["value"], d["config"]["sequential_ms_per_pair"], d["clocks"])
P
