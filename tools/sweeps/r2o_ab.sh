# in-call A/B of two builds of libspu_b200.so (tools/sweeps/libspu_b200_{A,B}.so), same box
mkdir -p gpurun_out
for rnd in 1 2; do
for v in A B; do
cp tools/sweeps/libspu_b200_$v.so starpu_inference_server_b200/libspu_b200.so
for c in C E; do
python bench.py --config $c --steps 50 --warmup 5 --no-cpu-baseline --no-offered-load > gpurun_out/r2o_${v}_${c}_$rnd.json 2> gpurun_out/r2o_${v}_${c}_$rnd.err || tail -3 gpurun_out/r2o_${v}_${c}_$rnd.err
done
python - <<PY
import json
for c in "CE":
    d=json.load(open("gpurun_out/r2o_${v}_%s_$rnd.json" % c)); print("$v", $rnd, c, round(d["value"]), round(d["value_sustained"]["value"]))
PY
done
done
