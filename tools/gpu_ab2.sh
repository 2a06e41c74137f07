#!/bin/bash
# bench-only A/B of experimental builds (PLTB200_LIB switch).  usage: gpu_ab2.sh "" _x _y ...
mkdir -p gpurun_out
for v in "$@"; do
  export PLTB200_LIB=mitsuba3_plt_b200/libpltb200$v.so
  [ -f $PLTB200_LIB ] || continue
  echo "== variant '$v'"
  python bench.py --no-cpu-baseline --steps 2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['kernel_ms_per_step'])"
done 2>&1 | tee gpurun_out/ab2.log
