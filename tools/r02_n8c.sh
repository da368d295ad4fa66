# Third 8-GPU call: bench C2 at N = 8 after bounding the parked tail to half of the pass (the SYRK of an 8-GPU shard no longer
# waits for the shard's whole upload).
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 3 --warmup 3 \
  > gpurun_out/r02n8c_bench_c2.json 2> gpurun_out/r02n8c_bench_c2.err; echo rc=$?; cat gpurun_out/r02n8c_bench_c2.json; tail -3 gpurun_out/r02n8c_bench_c2.err
