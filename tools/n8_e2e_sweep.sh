run() { name=$1; shift; timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 300 --warmup 10 "$@" > gpurun_out/s42_n8_$name.json 2> gpurun_out/s42_n8_$name.err; python -c "import sys,json; d=json.loads(open('gpurun_out/s42_n8_$name.json').read().strip().splitlines()[-1]); print('$name', round(d['value']), round(d['e2e']['value']), d['e2e']['how'], d['e2e']['task_latency_ms']['p50'])"; }
run autograph --auto-graph
SPU_BENCH_WORKERS=2 run w2
SPU_BENCH_WORKERS=2 SPU_BENCH_PIPELINE=4 run w2p4 --auto-graph
