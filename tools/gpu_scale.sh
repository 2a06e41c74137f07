#!/bin/bash
# strong-scaling check at N GPUs of one box: bench.py under torchrun (as the driver launches it)
N=${1:-8}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
tail -n 4 gpurun_out/scale_n$N.err; cat gpurun_out/scale_n$N.json
