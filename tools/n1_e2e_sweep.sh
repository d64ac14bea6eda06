run() { name=$1; shift; python bench.py --no-cpu-baseline --steps 300 "$@" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$name', round(d['value']), round(d['e2e']['value']), d['e2e']['how'], d['e2e']['task_latency_ms']['p50'])"; }
run w4p2_graph --auto-graph
SPU_BENCH_WORKERS=2 SPU_BENCH_PIPELINE=4 run w2p4_graph --auto-graph
run w4p2_eager
SPU_BENCH_WORKERS=2 SPU_BENCH_PIPELINE=4 run w2p4_eager
SPU_BENCH_WORKERS=3 SPU_BENCH_PIPELINE=3 run w3p3_graph --auto-graph
