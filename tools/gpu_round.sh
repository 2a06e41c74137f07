#!/bin/bash
# One GPU-box round: gpu tests, both bench arms, ncu launch list and a full-set capture of one pass (text pages exported
# on the box; the .ncu-rep itself is dropped when it would push gpurun_out/ over the 64 MiB pull limit).
T=${1:-r1}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/${T}_tests.log
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
python tools/ncu_target.py cbox-grating 1024 8 2 > gpurun_out/${T}_target.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python tools/ncu_target.py cbox-grating 1024 8 2 > gpurun_out/${T}_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -o /tmp/${T}_full -f \
    python tools/ncu_target.py cbox-grating 1024 8 1 > gpurun_out/${T}_ncu_f.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page details > gpurun_out/${T}_details.txt 2>/dev/null
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_raw.csv 2>/dev/null
ncu -i /tmp/${T}_full.ncu-rep --page source --csv --kernel-name regex:"k_shade|k_nee|k_trace" > gpurun_out/${T}_source.csv 2>/dev/null
gzip -f gpurun_out/${T}_source.csv gpurun_out/${T}_raw.csv
sz=$(stat -c %s /tmp/${T}_full.ncu-rep 2>/dev/null || echo 0)
if [ "$sz" -lt 40000000 ] && [ "$sz" -gt 0 ]; then cp /tmp/${T}_full.ncu-rep gpurun_out/; fi
cat gpurun_out/${T}_tests.log gpurun_out/${T}_bench.json
du -sh gpurun_out; ls -la gpurun_out
