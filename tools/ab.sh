# usage: tools/ab.sh "ENV=VAL ..." "ENV=VAL ..." ...   -- one short bench per variant, prints Gpx/s
mkdir -p gpurun_out; : > gpurun_out/ab.txt
i=0
for v in "$@"; do
  env $v timeout 300 python bench.py --steps 20 --no-e2e --no-cpu > gpurun_out/ab_$i.json 2>> gpurun_out/ab.err
  python - "$v" gpurun_out/ab_$i.json >> gpurun_out/ab.txt <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); print(f"{sys.argv[1]:50s} {d['value']/1e3:8.1f} Gpx/s  frac {d['roofline']['frac']:.4f}  {d['ms_per_step']:.3f} ms")
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
  i=$((i+1))
done
cat gpurun_out/ab.txt
