python bench.py --steps 6 --warmup 2 > gpurun_out/r02_bench_full1.json 2> gpurun_out/r02_bench_full1.err; echo rc=$?
tail -c 3000 gpurun_out/r02_bench_full1.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_full1.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","e2e","gpu_launches","launches_per_pair","d16","c5","c4","sharded","knn_tensor","cpu_baseline","clocks"):
    print(k, json.dumps(d.get(k))[:900])
print("roofline", {k:v for k,v in d["roofline"].items() if k not in ("note","effective_gbs_note")})
print("classes", d["kernel_classes"])
P
