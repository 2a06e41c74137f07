#!/bin/bash
# GPU-box round: gpu tests, gratingcap parity, C5 animation bench
T=${1:-s2}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/${T}_tests.log
python tools/parity_check.py gratingcap 256 16 0 2>&1 | tail -1 >> gpurun_out/${T}_parity.jsonl
python tools/anim_bench.py > gpurun_out/${T}_anim.json 2> gpurun_out/${T}_anim.err
cat gpurun_out/${T}_tests.log gpurun_out/${T}_parity.jsonl gpurun_out/${T}_anim.json
tail -n 3 gpurun_out/${T}_anim.err
