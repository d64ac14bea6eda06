# final record of the round: every BASELINE config through bench.py + the reference arm (one box)
mkdir -p gpurun_out
for c in B A C D E; do
  python bench.py --config $c --steps 200 --warmup 10 > gpurun_out/r2_bench_$c.json 2> gpurun_out/r2_bench_$c.err || { echo "config $c failed"; tail -5 gpurun_out/r2_bench_$c.err; }
done
python bench.py --impl reference --config B --steps 20 --warmup 5 > gpurun_out/r2_bench_reference_B.json 2> gpurun_out/r2_ref_B.err
python bench.py --config B --steps 20 --warmup 5 > gpurun_out/r2_bench_B_driver_args.json 2> gpurun_out/r2_bench_B_driver_args.err
python - <<'PY'
import json
for c in ["B","A","C","D","E","B_driver_args"]:
    try:
        d = json.load(open(f"gpurun_out/r2_bench_{c}.json"))
        p = d.get("parity") or {}
        print(c, round(d["value"]), "sust", round(d["value_sustained"]["value"]), "e2e", round(d["e2e"]["value"]), "deliv", round(d["e2e"]["delivered"]["value"]),
              "dv", d["e2e"]["device_visible"] and round(d["e2e"]["device_visible"]["value"]),
              "roof", round(d["roofline"]["frac"], 3), d["roofline"].get("kernel_ms_per_step"), "par", p.get("max_abs_over_max_ref"), p.get("top1_strict"), p.get("top1_margin_1e-2"),
              "cpu", round(d.get("cpu_baseline", {}).get("value", 0), 1), d.get("cpu_baseline", {}).get("kind"), "clk", d["clocks"]["samples"], d["clocks"]["reasons"])
    except Exception as e:
        print(c, "failed", e)
d = json.load(open("gpurun_out/r2_bench_reference_B.json")); print("reference", d["value"], d["steps"], d["warmup"], d["cpu_baseline"]["kind"], d["cpu_baseline"]["b1"])
PY
