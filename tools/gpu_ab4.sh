#!/bin/bash
# A/B of the block-rebalanced lobe scan of k_nee_post<grating> and of the BVH leaf size; every command under its own timeout
mkdir -p gpurun_out
for v in "" _rb _leaf1 _leaf2 _leaf8; do
  export PLTB200_LIB=mitsuba3_plt_b200/libpltb200$v.so
  [ -f $PLTB200_LIB ] || continue
  echo "== variant '$v'"
  timeout 150 python bench.py --no-cpu-baseline --steps 2 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['kernel_ms_per_step'], {k: round(v['ms'], 1) for k, v in d['kernels'].items()})"
done 2>&1 | tee gpurun_out/ab4.log
PLTB200_LIB=mitsuba3_plt_b200/libpltb200_rb.so timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/ab4.log
unset PLTB200_LIB
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/ab4.log
