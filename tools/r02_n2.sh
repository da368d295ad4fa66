# 2-GPU call of round 2 (gpurun --gpus 2 --timeout 900 -- 'bash tools/r02_n2.sh'): the multi-rank code of this round on real peers
# before it goes to the 8-GPU box: piece-grain band-pipelined reduce-scatter (bench + C++ engine), bench C4 chromosome shards,
# the CLI on two GPUs (SNP shards; LOCO chromosome shards with the output preparer).
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02n2_topo.txt 2>&1
timeout 300 python -m pytest tests/test_multiproc_gpu.py tests/test_host_engine.py -m gpu -q -x -k "multi or shard or tile or relay or pipelined or reduce or fetch_block or gapped or split_k" > gpurun_out/r02n2_tests.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02n2_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 2 --warmup 2 \
  > gpurun_out/r02n2_bench_c2.json 2> gpurun_out/r02n2_bench_c2.err; echo rc=$?; cat gpurun_out/r02n2_bench_c2.json; tail -3 gpurun_out/r02n2_bench_c2.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --config c4 --wall --gpus 2 --steps 2 --warmup 1 \
  > gpurun_out/r02n2_bench_c4.json 2> gpurun_out/r02n2_bench_c4.err; echo rc=$?; cat gpurun_out/r02n2_bench_c4.json; tail -3 gpurun_out/r02n2_bench_c4.err
