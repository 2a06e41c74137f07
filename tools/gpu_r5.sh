#!/bin/bash
# tests + bench (quick)
T=${1:-s5}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/${T}_tests.log
python tools/parity_check.py cbox-grating 256 16 0 2>&1 | tail -1 > gpurun_out/${T}_parity.jsonl
python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python tools/anim_bench.py > gpurun_out/${T}_anim.json 2>> gpurun_out/${T}_bench.err
cat gpurun_out/${T}_tests.log gpurun_out/${T}_parity.jsonl
python - <<PY
import json
d = json.load(open("gpurun_out/${T}_bench.json")); print(d["value"], d["e2e"]["value"], d["kernel_ms_per_step"])
PY
cat gpurun_out/${T}_anim.json; tail -n 5 gpurun_out/${T}_bench.err
