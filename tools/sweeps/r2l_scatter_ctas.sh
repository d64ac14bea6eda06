mkdir -p gpurun_out
for c in 1024 64 32 128; do
SPU_SCATTER_CTAS=$c python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-offered-load > gpurun_out/r2l_bench_$c.json 2>gpurun_out/r2l_bench_$c.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2l_bench_$c.json")); print($c, round(d["value"]), round(d["e2e"]["value"]), round(d["e2e"]["delivered"]["value"]), round(d["e2e"]["device_visible"]["value"]))
PY
done
