#!/bin/bash
# r01w: C5 in 2D-tile mode on N GPUs, BED resident on every GPU
N=${1:-8}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tools/tile_probe.py > gpurun_out/r01w_tile_n$N.json 2> gpurun_out/r01w_tile_n$N.err
echo "rc=$?"; cat gpurun_out/r01w_tile_n$N.json; tail -5 gpurun_out/r01w_tile_n$N.err
