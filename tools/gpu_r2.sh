#!/bin/bash
# GPU-box round: gpu tests, parity on all config scenes, bench (both arms), C5 animation bench, launch list.
T=${1:-s1}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_tests.log
for sc in cbox-grating veach-mis cbox gratingcap; do
  python tools/parity_check.py $sc 256 16 0 2>&1 | tail -1 >> gpurun_out/${T}_parity.jsonl
done
python tools/anim_bench.py > gpurun_out/${T}_anim.json 2> gpurun_out/${T}_anim.err
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python tools/ncu_target.py cbox-grating 1024 8 2 > gpurun_out/${T}_ncu_l.log 2>&1
cat gpurun_out/${T}_tests.log gpurun_out/${T}_parity.jsonl gpurun_out/${T}_anim.json gpurun_out/${T}_bench.json gpurun_out/${T}_bench_ref.json
tail -3 gpurun_out/${T}_anim.err gpurun_out/${T}_bench.err
