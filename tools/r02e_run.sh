# Fifth GPU call (1 GPU): the four tests that failed in r02d, the host-register probe, C2 CLI after the multi-GPU engine edits.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x -s -k "fetch_block or gapped or split_k or tile_mode_large or engine_multi_gpu" > gpurun_out/r02e_pytest.log 2>&1; echo rc=$?; tail -5 gpurun_out/r02e_pytest.log
timeout 120 python tools/probes/hostreg_probe.py /dev/shm/ggrm_hostreg_probe 4 > gpurun_out/r02e_hostreg.json 2>&1; echo rc=$?; cat gpurun_out/r02e_hostreg.json
timeout 120 python tools/probes/hostreg_probe.py /tmp/ggrm_hostreg_probe 2 > gpurun_out/r02e_hostreg_tmp.json 2>&1; echo rc=$?; cat gpurun_out/r02e_hostreg_tmp.json
GELEX_B200_OVERSUBSCRIBE=1 timeout 200 python tools/cli_configs.py C2 --precision bf16 > gpurun_out/r02e_c2.jsonl 2> gpurun_out/r02e_c2.err; echo rc=$?; cat gpurun_out/r02e_c2.jsonl
