# Round-1 run D: GPU tests, all named configs on one GPU (HBM-resident BED), box facts.
set -x
nproc; free -g | head -2; df -h /tmp /dev/shm . | cat
python -m pytest tests -m gpu -x -q > gpurun_out/r01d_pytest.log 2>&1; tail -3 gpurun_out/r01d_pytest.log
python tools/config_probe.py C1 C3 C2 C4 --precision bf16 > gpurun_out/r01d_configs_bf16.jsonl 2> gpurun_out/r01d_configs.err; echo rc=$?
python tools/config_probe.py C5 --precision bf16 --passes 1 >> gpurun_out/r01d_configs_bf16.jsonl 2>> gpurun_out/r01d_configs.err; echo rc=$?
python tools/config_probe.py C3 C4 --precision fp16 > gpurun_out/r01d_configs_fp16.jsonl 2>> gpurun_out/r01d_configs.err; echo rc=$?
cat gpurun_out/r01d_configs_bf16.jsonl gpurun_out/r01d_configs_fp16.jsonl
