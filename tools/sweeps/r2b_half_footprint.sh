set -x
python -m pytest tests/test_gpu_stages.py -x -q -k "half_footprint or every_tile or tcgen05_matches or inplace" > gpurun_out/r2b_tests.log 2>&1; echo "exit $?" >> gpurun_out/r2b_tests.log
tail -5 gpurun_out/r2b_tests.log
for st in 1 2; do
python tools/gemm_bench.py --streams $st --shapes attn_out,ffn2 --cfgs 2:192,3:192,3:128,2:128 >> gpurun_out/r2b_gemm.jsonl 2>&1
python tools/gemm_bench.py --streams $st --shapes qkv,ffn1 --cfgs 2:256,3:256,3:192,3:128,2:128 >> gpurun_out/r2b_gemm.jsonl 2>&1
done
cat gpurun_out/r2b_gemm.jsonl
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r2b_bench_full.json 2>gpurun_out/r2b_bench_full.err
SPU_GEMM_HF=1 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r2b_bench_hf.json 2>gpurun_out/r2b_bench_hf.err
python - <<'PY'
import json
for n in ("full","hf"):
    try:
        d=json.load(open(f"gpurun_out/r2b_bench_{n}.json"))
        print(n, d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms_per_step"], d["stage_ms_per_step"])
    except Exception as e: print(n, "failed", e)
PY
