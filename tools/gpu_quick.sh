#!/bin/bash
# Quick GPU-box round: gpu tests, large-sample parity, bench (no cpu baseline), launch list, SASS-level source page.
T=${1:-q}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_tests.log
python tools/parity_check.py cbox-grating 256 16 0 > gpurun_out/${T}_parity.json 2>&1
python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python tools/ncu_target.py cbox-grating 1024 8 2 > gpurun_out/${T}_ncu_l.log 2>&1
if [ "$2" = "full" ]; then
ncu --set full --clock-control none --import-source on -o /tmp/${T}_full -f \
    python tools/ncu_target.py cbox-grating 1024 8 1 > gpurun_out/${T}_ncu_f.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page details > gpurun_out/${T}_details.txt 2>/dev/null
ncu -i /tmp/${T}_full.ncu-rep --page source --csv > gpurun_out/${T}_source.csv 2>/dev/null
gzip -f gpurun_out/${T}_source.csv
fi
cat gpurun_out/${T}_tests.log gpurun_out/${T}_parity.json gpurun_out/${T}_bench.json
du -sh gpurun_out
