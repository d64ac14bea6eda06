# A/B of the completion-wait mode on the e2e path (prints value, e2e, segments)
for mode in block spin block spin; do
  if [ $mode = spin ]; then export SPU_SPIN_SYNC=1; else unset SPU_SPIN_SYNC; fi
  python bench.py --no-cpu-baseline --steps 300 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$mode', round(d['value']), round(d['e2e']['value']), d['e2e']['how'], d['e2e']['task_latency_ms']['p50'])"
done
