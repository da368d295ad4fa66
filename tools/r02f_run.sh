# Sixth GPU call (1 GPU): direct DMA into page-locked output mappings - byte test, then C2 / C3 / C4 CLI walls with it on and off.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_host_engine.py -m gpu -q -x > gpurun_out/r02f_pytest.log 2>&1; echo rc=$?; tail -5 gpurun_out/r02f_pytest.log
for dma in 1 0; do
  GELEX_B200_DIRECT_DMA=$dma timeout 400 python tools/cli_configs.py C2 C3 C4 --precision fp16 > gpurun_out/r02f_cli_dma$dma.jsonl 2> gpurun_out/r02f_cli_dma$dma.err; echo rc=$?; cat gpurun_out/r02f_cli_dma$dma.jsonl
done
