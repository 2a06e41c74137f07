#!/bin/bash
# tests + parity + bench A/B of the strip splat against the direct per-tap atomics
T=${1:-s4}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/${T}_tests.log
for sc in cbox-grating veach-mis spectrograph:d65; do
  python tools/parity_check.py $sc 256 16 0 2>&1 | tail -1 >> gpurun_out/${T}_parity.jsonl
done
python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
PLTB_SPLAT_DIRECT=1 python bench.py --no-cpu-baseline --steps 2 > gpurun_out/${T}_bench_direct.json 2>> gpurun_out/${T}_bench.err
python tools/anim_bench.py > gpurun_out/${T}_anim.json 2>> gpurun_out/${T}_bench.err
cat gpurun_out/${T}_tests.log gpurun_out/${T}_parity.jsonl
python - <<PY
import json
for f in ("gpurun_out/${T}_bench.json", "gpurun_out/${T}_bench_direct.json"):
    d = json.load(open(f)); print(f, d["value"], d["e2e"]["value"], d["kernel_ms_per_step"])
PY
cat gpurun_out/${T}_anim.json; tail -n 5 gpurun_out/${T}_bench.err
