#!/bin/bash
# Round-1 head (4fdeb6e, checked out and built under _ab/r01 by `git worktree add _ab/r01 4fdeb6e` +
# `python -m starpu_inference_server_b200.build` there) against this tree, interleaved A B A B on the box this runs on.
out=gpurun_out/r3_bisect.jsonl
: > $out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader >> gpurun_out/r3_bisect_box.txt
for i in 1 2; do
  python tools/sweeps/r3_ab_chain.py /root/repo/_ab/r01 r01 >> $out 2>> gpurun_out/r3_bisect.err
  python tools/sweeps/r3_ab_chain.py /root/repo HEAD >> $out 2>> gpurun_out/r3_bisect.err
done
(cd _ab/r01 && python bench.py --steps 200 --warmup 5 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps({'label':'r01 bench','value':d['value'],'ms_per_step':d['ms_per_step'],'e2e':d['e2e']['value'],'roofline':d['roofline'].get('kernel_ms_per_step'),'clocks':d.get('clocks')}))") >> $out
python bench.py --steps 200 --warmup 5 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps({'label':'HEAD bench','value':d['value'],'ms_per_step':d['ms_per_step'],'e2e':d['e2e']['value'],'roofline':d['roofline'].get('kernel_ms_per_step'),'clocks':d.get('clocks')}))" >> $out
cat $out
