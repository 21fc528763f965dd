python -m pytest tests -m gpu -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_final_1gpu.json 2> gpurun_out/r02_bench_final_1gpu.err; echo rc=$?
tail -c 400 gpurun_out/r02_bench_final_1gpu.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_final_1gpu.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","e2e","gpu_launches","launches_per_pair","d16","c5","c4","sharded","knn_tensor","cpu_baseline","clocks"):
    print(k, json.dumps(d.get(k))[:700])
print("roofline", {k:v for k,v in d["roofline"].items() if k not in ("note","effective_gbs_note")})
P
