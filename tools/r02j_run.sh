set -x
mkdir -p gpurun_out
timeout 120 python tools/config_probe.py C1 --passes 6 --precision fp16 > gpurun_out/r02j_c1.jsonl 2>&1; echo rc=$?; cat gpurun_out/r02j_c1.jsonl
timeout 300 python -m pytest tests -m gpu -q -x -k "split_k or c1_ or fetch_block or tile_mode or 16bit_vs_oracle or guard or smoke or no_alloc" > gpurun_out/r02j_pytest.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02j_pytest.log
