# 2-GPU regression of the final code on real peers: the multi-rank tests (GCTA diagonal reduce over peer memory included) and one bench line.
set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gcta_diag_gpu.py tests/test_multiproc_gpu.py tests/test_host_engine.py -m gpu -q -x -k "engine_gcta or multi or shard or tile or relay or pipelined or reduce or piece" > gpurun_out/r02zf_tests.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02zf_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 2 --warmup 3 --no-wall \
  > gpurun_out/r02zf_bench_c2_n2.json 2> gpurun_out/r02zf_bench_c2_n2.err; echo rc=$?; cut -c1-900 gpurun_out/r02zf_bench_c2_n2.json; tail -3 gpurun_out/r02zf_bench_c2_n2.err
