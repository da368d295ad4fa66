# Second 8-GPU call of round 2: bench C2 / C4 at N = 8 after the short first slab, the final-strip band, lazy writer pools and the
# smaller pinned ring; both with their CLI wall legs.
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 3 --warmup 3 \
  > gpurun_out/r02n8b_bench_c2.json 2> gpurun_out/r02n8b_bench_c2.err; echo rc=$?; cat gpurun_out/r02n8b_bench_c2.json; tail -3 gpurun_out/r02n8b_bench_c2.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --config c4 --wall --gpus 8 --steps 2 --warmup 2 \
  > gpurun_out/r02n8b_bench_c4.json 2> gpurun_out/r02n8b_bench_c4.err; echo rc=$?; cat gpurun_out/r02n8b_bench_c4.json; tail -3 gpurun_out/r02n8b_bench_c4.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 3 --warmup 3 --no-wall \
  > gpurun_out/r02n8b_bench_c2_n4.json 2> gpurun_out/r02n8b_bench_c2_n4.err; echo rc=$?; cat gpurun_out/r02n8b_bench_c2_n4.json
