#!/bin/bash
# A/B of the k_nee_post<grating> variants (block-rebalanced lobe scan, register caps): bench, then the GPU parity suite on the
# rebalanced build
bash tools/gpu_ab2.sh "" _rb _mb6 _rb6
PLTB200_LIB=mitsuba3_plt_b200/libpltb200_rb.so python -m pytest tests -m gpu -x -q 2>&1 | tail -3
