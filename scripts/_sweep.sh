for c in 1 2 3 4 6; do
  CB_KNN_HALFWAVES=$c ncu --metrics gpu__time_duration.sum --clock-control none -k regex:knn_ -c 12 --csv --log-file gpurun_out/kn_$c.csv python scripts/ransac_probe.py 4096 2 1 > /dev/null 2>&1
  echo "halfwaves $c:"; python - <<PY
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/kn_$c.csv") if l.startswith('"'))]
for r in rows[1:7]: print("   ", r[4][:40], r[8], r[-1])
PY
done
