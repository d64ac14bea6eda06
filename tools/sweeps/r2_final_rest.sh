mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gputests_final.log 2>&1; tail -2 gpurun_out/r2_gputests_final.log
for c in A C D E; do
  python bench.py --config $c --steps 200 --warmup 10 > gpurun_out/r2_bench_$c.json 2> gpurun_out/r2_bench_$c.err || { echo "config $c failed"; tail -5 gpurun_out/r2_bench_$c.err; }
done
python bench.py --config B --steps 200 --warmup 10 --no-offered-load > gpurun_out/r2_bench_B_thisbox.json 2> gpurun_out/r2_bench_B_thisbox.err
python - <<'PY'
import json
for c in ["A","C","D","E","B_thisbox"]:
    d = json.load(open(f"gpurun_out/r2_bench_{c}.json")); p = d.get("parity") or {}; r = d["roofline"]
    print(c, round(d["value"]), "sust", round(d["value_sustained"]["value"]), "e2e", round(d["e2e"]["value"]), "deliv", round(d["e2e"]["delivered"]["value"]),
          "dv", d["e2e"]["device_visible"] and round(d["e2e"]["device_visible"]["value"]), "roof", round(r["frac"], 3), r.get("frac_of_sustained_peak"), r.get("kernel_ms_per_step"),
          "par", p.get("max_abs_over_max_ref"), p.get("top1_strict"), p.get("top1_margin_1e-2"), "cpu", round(d.get("cpu_baseline", {}).get("value", 0), 1), "d2h", d["e2e"]["d2h_bytes_per_step"])
    for l in (d.get("offered_load") or {}).get("loads", []): print("    ", l["offered_fraction_of_e2e"], round(l["achieved_per_s"]), round(l["mean_batch_rows"],1), l["task_ms"], l["request_ms"])
PY
