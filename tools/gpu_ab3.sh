#!/bin/bash
# smoke + A/B of splat sample-chunk variants and of the persistent-grid multiplier
python __graft_entry__.py smoke 2>&1 | tail -2
bash tools/gpu_ab2.sh "" _ss16 _ss32
for g in 4 16 32; do echo "grid mult $g"; PLTB_GRID_MULT=$g python bench.py --no-cpu-baseline --steps 2 | python -c "import sys,json; d=json.load(sys.stdin); print(d['value'], d['kernel_ms_per_step'])"; done
