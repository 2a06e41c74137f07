#!/bin/bash
# confirmation of the committed build: gpu tests, smoke, both bench arms
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/v11_tests.log
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -1 | cut -c1-160
timeout 400 python bench.py > gpurun_out/v11_bench.json 2> gpurun_out/v11_bench.err; cut -c1-400 gpurun_out/v11_bench.json
