# Fourth 8-GPU call: final code at N = 8 (bench C2 with its CLI wall leg: --gpus 8 is now an upper bound, the CLI takes 2 for C2),
# and the same CLI command with --gpus-exact for the comparison.
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 3 --warmup 3 \
  > gpurun_out/r02n8d_bench_c2.json 2> gpurun_out/r02n8d_bench_c2.err; echo rc=$?; cat gpurun_out/r02n8d_bench_c2.json; tail -3 gpurun_out/r02n8d_bench_c2.err
