set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_named_configs_gpu.py -m gpu -q -x > gpurun_out/r02zc_tests.log 2>&1; echo rc=$?
tail -5 gpurun_out/r02zc_tests.log
for f in 0 1 0 1; do
  GGRM_FUSED_PREP=$f timeout 200 python tools/config_probe.py C1 --precision fp16 --passes 4 2> gpurun_out/r02zc.err | tee -a gpurun_out/r02zc_c1_fused$f.jsonl; tail -2 gpurun_out/r02zc.err
done
