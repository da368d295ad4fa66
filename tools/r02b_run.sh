# Second GPU call of round 2 (1 GPU): the round's library changes (split-K, BF16 hi+lo, FP16 overflow flag, parked-slab bound), the
# named-config parity tests, both bench configs at N = 1, and C1 with / without split-K.
#   gpurun --timeout 1700 -- 'bash tools/r02b_run.sh'
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "loco_and_normalised or gapped or split_k or hi_lo or overflow or 16bit_vs_oracle or c1_shape" > gpurun_out/r02b_new_tests.log 2>&1; echo rc=$?; tail -15 gpurun_out/r02b_new_tests.log
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r02b_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r02b_smoke.log
timeout 900 python -m pytest tests/test_named_configs_gpu.py -m gpu -q -s > gpurun_out/r02b_named.log 2>&1; echo rc=$?; grep -E "max|passed|failed|Error" gpurun_out/r02b_named.log | tail -30
for ks in 1 0; do
  if [ $ks = 1 ]; then export GGRM_KSPLIT=1; else unset GGRM_KSPLIT; fi
  timeout 120 python tools/config_probe.py C1 --passes 4 > gpurun_out/r02b_c1_ksplit_$ks.jsonl 2>&1; echo rc=$?; cat gpurun_out/r02b_c1_ksplit_$ks.jsonl
done
unset GGRM_KSPLIT
timeout 400 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r02b_bench_c2.json 2> gpurun_out/r02b_bench_c2.err; echo rc=$?; cat gpurun_out/r02b_bench_c2.json; tail -5 gpurun_out/r02b_bench_c2.err
timeout 400 python bench.py --config c4 --gpus 1 --steps 2 --warmup 2 --no-cpu > gpurun_out/r02b_bench_c4.json 2> gpurun_out/r02b_bench_c4.err; echo rc=$?; cat gpurun_out/r02b_bench_c4.json; tail -5 gpurun_out/r02b_bench_c4.err
timeout 600 python -m pytest tests -m gpu -q -x --deselect tests/test_named_configs_gpu.py > gpurun_out/r02b_pytest_rest.log 2>&1; echo rc=$?; tail -8 gpurun_out/r02b_pytest_rest.log
