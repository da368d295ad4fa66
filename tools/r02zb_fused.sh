# fused single-slab prep: bit-identity test + the parity tests that now run through it, then C1 with and without it
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_named_configs_gpu.py tests/test_gcta_diag_gpu.py -m gpu -q -x > gpurun_out/r02zb_tests.log 2>&1; echo rc=$?
tail -25 gpurun_out/r02zb_tests.log
for f in 0 1; do
  GGRM_FUSED_PREP=$f timeout 200 python tools/config_probe.py C1 --precision fp16 --passes 4 > gpurun_out/r02zb_c1_fused$f.jsonl 2> gpurun_out/r02zb_c1_fused$f.err; echo rc=$?
  cat gpurun_out/r02zb_c1_fused$f.jsonl; tail -3 gpurun_out/r02zb_c1_fused$f.err
done
GGRM_FUSED_PREP=1 timeout 200 python tools/config_probe.py C1 --precision bf16 --passes 4 > gpurun_out/r02zb_c1_bf16.jsonl 2>&1; cat gpurun_out/r02zb_c1_bf16.jsonl
