# 1/2/4/8-GPU record on ONE box: torchrun bench (one process per GPU) for config B, N=8 for C, D, E, and the
# one-process / one-scheduler topology (tools/one_process_scaling.py)
mkdir -p gpurun_out
run() {  # n config
  if [ "$1" = 1 ]; then python bench.py --config $2 --steps 200 --warmup 10 --no-cpu-baseline --no-offered-load > gpurun_out/r2_scale_$2_n$1.json 2> gpurun_out/r2_scale_$2_n$1.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 2960$1 bench.py --config $2 --gpus $1 --steps 200 --warmup 10 > gpurun_out/r2_scale_$2_n$1.json 2> gpurun_out/r2_scale_$2_n$1.err; fi
}
for n in 1 2 4 8; do run $n B; done
for c in C D E; do run 1 $c; run 8 $c; done
python tools/one_process_scaling.py --steps 300 > gpurun_out/r2_one_process.jsonl 2> gpurun_out/r2_one_process.err
python - <<'PY'
import json
base = {}
for c in "BCDE":
    for n in (1, 2, 4, 8):
        try:
            d = json.load(open(f"gpurun_out/r2_scale_{c}_n{n}.json"))
        except Exception as e:
            continue
        v, s, e, dl = d["value"], d["value_sustained"]["value"], d["e2e"]["value"], d["e2e"]["delivered"]["value"]
        if n == 1: base[c] = (v, s, e, dl)
        b = base.get(c, (v, s, e, dl))
        print(c, n, round(v), f"{v/b[0]:.2f}x", "sust", round(s), f"{s/b[1]:.2f}x", "e2e", round(e), f"{e/b[2]:.2f}x", "deliv", round(dl), f"{dl/b[3]:.2f}x", d["run"]["cores_per_rank"])
print(open("gpurun_out/r2_one_process.jsonl").read())
PY
