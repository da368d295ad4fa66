# 8-GPU call of round 2 (gpurun --gpus 8 --timeout 900 -- 'bash tools/r02_n8.sh'): what has never been measured on a full box.
#   1. multi-process / multi-GPU tests on real peers (NVLink P2P instead of two ranks sharing one device)
#   2. bench C2 at N = 8 under torchrun, as the driver launches it (piece-grain band-pipelined reduce-scatter; CLI wall leg at 8 GPUs)
#   3. bench C4 at N = 8: 22 LOCO partials + whole, chromosome shards; CLI wall leg writes 23 x 5 GB
#   4. C5 (200k x 600k) through the CLI in 2D-tile mode with the BED relay
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02n8_topo.txt 2>&1; df -h /dev/shm /tmp > gpurun_out/r02n8_df.txt 2>&1; nproc >> gpurun_out/r02n8_df.txt; free -g >> gpurun_out/r02n8_df.txt
timeout 240 python -m pytest tests/test_multiproc_gpu.py tests/test_host_engine.py -m gpu -q -x -k "multi or shard or tile or relay or pipelined or reduce" > gpurun_out/r02n8_tests.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02n8_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 3 --warmup 3 \
  > gpurun_out/r02n8_bench_c2.json 2> gpurun_out/r02n8_bench_c2.err; echo rc=$?; cat gpurun_out/r02n8_bench_c2.json; tail -3 gpurun_out/r02n8_bench_c2.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --config c4 --wall --gpus 8 --steps 2 --warmup 2 \
  > gpurun_out/r02n8_bench_c4.json 2> gpurun_out/r02n8_bench_c4.err; echo rc=$?; cat gpurun_out/r02n8_bench_c4.json; tail -3 gpurun_out/r02n8_bench_c4.err
GELEX_B200_BED_RELAY=1 timeout 300 python tools/cli_configs.py C5 --gpus 8 > gpurun_out/r02n8_c5_relay.jsonl 2> gpurun_out/r02n8_c5_relay.err; echo rc=$?; cat gpurun_out/r02n8_c5_relay.jsonl; tail -3 gpurun_out/r02n8_c5_relay.err
