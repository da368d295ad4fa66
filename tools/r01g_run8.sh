# Round-1 run G (8 GPUs of one box): bench at N = 2, 4, 8 as the driver launches it, then C2 (SNP shards) and C5 (2D tiles)
# through the CLI with real files on tmpfs.
set -x
nvidia-smi -L | head -8; free -g | head -2; nproc
for N in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 3 --warmup 3 \
    > gpurun_out/r01g_bench_n$N.json 2> gpurun_out/r01g_bench_n$N.err; echo rc=$?; tail -c 2500 gpurun_out/r01g_bench_n$N.json
done
timeout 300 python tools/cli_configs.py C2 --gpus 8 --precision bf16 > gpurun_out/r01g_cli.jsonl 2> gpurun_out/r01g_cli.err; echo rc=$?
avail=$(free -g | awk '/Mem:/ {print $7}')
if [ "$avail" -gt 400 ]; then
  timeout 600 python tools/cli_configs.py C5 --gpus 8 --precision bf16 >> gpurun_out/r01g_cli.jsonl 2>> gpurun_out/r01g_cli.err; echo rc=$?
fi
cat gpurun_out/r01g_cli.jsonl; tail -5 gpurun_out/r01g_cli.err
