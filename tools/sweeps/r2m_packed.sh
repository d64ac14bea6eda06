mkdir -p gpurun_out
python -m pytest tests/test_gpu_forward.py tests/test_gpu_host.py tests/test_gpu_stages.py -x -q > gpurun_out/r2m_tests.log 2>&1; tail -3 gpurun_out/r2m_tests.log
for v in "" "--no-packed"; do
python bench.py --config C --steps 50 --warmup 5 $v > gpurun_out/r2m_bench_C$v.json 2> gpurun_out/r2m_bench_C$v.err || tail -5 gpurun_out/r2m_bench_C$v.err
done
python - <<'PY'
import json
for v in ("", "--no-packed"):
    try:
        d=json.load(open(f"gpurun_out/r2m_bench_C{v}.json")); p=d["parity"]
        print(v or "packed", round(d["value"]), round(d["value_sustained"]["value"]), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["delivered"]["value"]), d["e2e"]["device_visible"] and round(d["e2e"]["device_visible"]["value"]),
              "roof", round(d["roofline"]["frac"],3), d["roofline"]["kernel_ms_per_step"], "par", p["max_abs_over_max_ref"], p["top1_strict"], p["top1_margin_1e-2"], d["stage_ms_per_step"])
        if d.get("offered_load"): print("   ", json.dumps(d["offered_load"]["loads"]))
    except Exception as e: print(v, "failed", e)
PY
