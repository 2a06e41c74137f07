#!/bin/bash
# A/B of experimental builds of libpltb200 (PLTB200_LIB switch): parity + bench for each.  usage: gpu_ab.sh "" _x _y ...
mkdir -p gpurun_out
for v in "$@"; do
  export PLTB200_LIB=mitsuba3_plt_b200/libpltb200$v.so
  [ -f $PLTB200_LIB ] || continue
  echo "== variant '$v'"
  for sc in cbox-grating veach-mis cbox; do
  python tools/parity_check.py $sc 256 16 0 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['scene'], d['fraction_within_1e-4'], d['image_rel_mse'], d['nan_mask_mismatch'], d['nan_samples_gpu'], d['nan_samples_oracle'])"
  done
  python bench.py --no-cpu-baseline --steps 2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['kernel_ms_per_step'])"
done 2>&1 | tee gpurun_out/ab.log
