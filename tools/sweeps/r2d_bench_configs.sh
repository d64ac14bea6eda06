# every BASELINE config through bench.py (one JSON line each) + the reference arm for B
mkdir -p gpurun_out
for c in B A E C D; do
  python bench.py --config $c --steps ${STEPS:-50} --warmup 5 > gpurun_out/r2d_bench_$c.json 2> gpurun_out/r2d_bench_$c.err || { echo "config $c failed"; tail -5 gpurun_out/r2d_bench_$c.err; }
done
python bench.py --impl reference --config B --steps 10 --warmup 3 > gpurun_out/r2d_ref_B.json 2> gpurun_out/r2d_ref_B.err
python - <<'PY'
import json
for c in "BAECD":
    try:
        d = json.load(open(f"gpurun_out/r2d_bench_{c}.json"))
        p = d.get("parity") or {}
        print(c, round(d["value"]), "sust", round(d["value_sustained"]["value"]), "e2e", round(d["e2e"]["value"]), "deliv", round(d["e2e"]["delivered"]["value"]),
              "roof", round(d["roofline"]["frac"], 3), "par", p.get("max_abs_over_max_ref"), p.get("top1_strict"), p.get("top1_margin_1e-2"),
              "cpu", d.get("cpu_baseline", {}).get("value"), d.get("cpu_baseline", {}).get("kind"), "clk", d["clocks"]["samples"])
        if d.get("offered_load"): print("   offered", json.dumps(d["offered_load"]["loads"]))
    except Exception as e:
        print(c, "failed", e)
print(open("gpurun_out/r2d_ref_B.json").read()[:600])
PY
