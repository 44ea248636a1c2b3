#!/bin/bash
# Runs on the GPU box (through gpurun): GPU test suite, then a short bench; prints the headline numbers.
# usage: tools/gpu_check.sh <tag> [pytest args...]
tag=${1:-check}; shift
python -m pytest tests -m gpu -x -q "$@" 2>&1 | tail -15
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/${tag}.json 2> gpurun_out/${tag}.err || tail -c 1500 gpurun_out/${tag}.err
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${tag}.json"))
    print({k: d[k] for k in ["value", "ms_per_step", "ms_per_iteration"]}, "e2e", d["e2e"]["value"])
    print({k: round(v * 1e3, 2) for k, v in d["roofline"]["stage_ms"].items()})
except Exception as e:
    print("bench line unreadable:", e)
PY
