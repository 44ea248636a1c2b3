#!/bin/bash
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs 2>gpurun_out/bench_only.err | python -c "
import json,sys
t=sys.stdin.read()
try:
    d=json.loads(t); print(round(d['value'],1), round(d['ms_per_iteration']*1e3,1), {k: round(v*1e3,2) for k,v in d['roofline']['stage_ms'].items()})
except Exception as e:
    print('no line', e); print(open('gpurun_out/bench_only.err').read()[-800:])"
