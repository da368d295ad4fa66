# 8-GPU call of the next round (gpurun --gpus 8 --timeout 1500 -- 'bash tools/r02_n8.sh'): what round 1 left unmeasured on a full box.
#   1. bench at N = 8 under torchrun (as the driver launches it)
#   2. C2 through the CLI on 8 GPUs (SNP shards) with both writers: round 1 had 7.4 s through pwrite (eight writers on one tmpfs inode)
#   3. C5 through the CLI (2D tiles) with pwrite / mapping, then with the BED relay (accumulate 5.1 s -> ~2.3 s expected)
#   4. C4 (22 LOCO partials + whole) on 8 GPUs, chromosome shards: the north star's "plus 22 LOCO GRMs" on a full box
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 3 --warmup 3 \
  > gpurun_out/r02n8_bench.json 2> gpurun_out/r02n8_bench.err; echo rc=$?; cat gpurun_out/r02n8_bench.json
for w in pwrite mmap; do
  GELEX_B200_WRITER=$w timeout 200 python tools/cli_configs.py C2 --gpus 8 > gpurun_out/r02n8_c2_$w.jsonl 2> gpurun_out/r02n8_c2_$w.err; echo rc=$?; cat gpurun_out/r02n8_c2_$w.jsonl
done
GELEX_B200_WRITER=mmap GELEX_B200_PREFAULT=1 timeout 200 python tools/cli_configs.py C2 --gpus 8 > gpurun_out/r02n8_c2_prefault.jsonl 2> gpurun_out/r02n8_c2_prefault.err; echo rc=$?; cat gpurun_out/r02n8_c2_prefault.jsonl
GELEX_B200_WRITER=pwrite timeout 300 python tools/cli_configs.py C5 --gpus 8 > gpurun_out/r02n8_c5_pwrite.jsonl 2> gpurun_out/r02n8_c5_pwrite.err; echo rc=$?; cat gpurun_out/r02n8_c5_pwrite.jsonl
GELEX_B200_WRITER=mmap timeout 300 python tools/cli_configs.py C5 --gpus 8 > gpurun_out/r02n8_c5_mmap.jsonl 2> gpurun_out/r02n8_c5_mmap.err; echo rc=$?; cat gpurun_out/r02n8_c5_mmap.jsonl
GELEX_B200_WRITER=mmap GELEX_B200_PREFAULT=1 GELEX_B200_BED_RELAY=1 timeout 300 python tools/cli_configs.py C5 --gpus 8 > gpurun_out/r02n8_c5_relay.jsonl 2> gpurun_out/r02n8_c5_relay.err; echo rc=$?; cat gpurun_out/r02n8_c5_relay.jsonl
timeout 400 python tools/cli_configs.py C4 --gpus 8 > gpurun_out/r02n8_c4.jsonl 2> gpurun_out/r02n8_c4.err; echo rc=$?; cat gpurun_out/r02n8_c4.jsonl
