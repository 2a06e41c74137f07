#!/bin/bash
# ncu pass: launch list of two renders + full-set capture of one pass (1024^2 x 8 spp); details + source pages exported.
T=${1:-p1}
mkdir -p gpurun_out
python tools/ncu_target.py cbox-grating 1024 8 2 > gpurun_out/${T}_target.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python tools/ncu_target.py cbox-grating 1024 8 2 > gpurun_out/${T}_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -o /tmp/${T}_full -f \
    python tools/ncu_target.py cbox-grating 1024 8 1 > gpurun_out/${T}_ncu_f.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page details > gpurun_out/${T}_details.txt 2>/dev/null
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_raw.csv 2>/dev/null
ncu -i /tmp/${T}_full.ncu-rep --page source --csv > gpurun_out/${T}_source.csv 2>/dev/null
gzip -f gpurun_out/${T}_source.csv gpurun_out/${T}_raw.csv
du -sh gpurun_out; tail -n 3 gpurun_out/${T}_ncu_f.log
