# Round-1 run I (8 GPUs): the GPU test-suite on real peers (IPC reduce across NVLink, engine decompositions on distinct
# devices), then bench N = 8 and 4 with ranks bound to their GPU's NUMA node.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01i_pytest.log 2>&1; tail -4 gpurun_out/r01i_pytest.log
for N in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) bench.py --gpus $N --steps 3 --warmup 3 \
    > gpurun_out/r01i_bench_n$N.json 2> gpurun_out/r01i_bench_n$N.err; echo rc=$?; tail -c 1800 gpurun_out/r01i_bench_n$N.json
done
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 bench.py --impl reference --gpus 8 --steps 1 --warmup 0 --cpu-budget 4 > gpurun_out/r01i_ref_n8.json 2> gpurun_out/r01i_ref_n8.err; echo rc=$?; tail -c 600 gpurun_out/r01i_ref_n8.json
