for hf in 0 2 3; do
SPU_GEMM_HF=$hf python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r2c_bench_hf$hf.json 2>gpurun_out/r2c_bench_hf$hf.err
done
python - <<'PY'
import json
for n in (0,2,3):
    try:
        d=json.load(open(f"gpurun_out/r2c_bench_hf{n}.json"))
        print(n, d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms_per_step"], d["stage_ms_per_step"])
    except Exception as e: print(n, "failed", e)
PY
