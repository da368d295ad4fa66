# Fourth GPU call of round 2 (1 GPU): whole GPU suite after the fixes of r02c; C1 through the probe with FP16 operands (unsplit);
# C4 through the CLI with 4 and 8 writer / preparer threads; pack_rows8 with column chunks (launch list).
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -s > gpurun_out/r02d_pytest.log 2>&1; echo rc=$?; grep -E "max abs|max \||unsplit|passed|failed|Error" gpurun_out/r02d_pytest.log | tail -60
for ks in 1 0; do
  if [ $ks = 1 ]; then export GGRM_KSPLIT=1; else unset GGRM_KSPLIT; fi
  timeout 120 python tools/config_probe.py C1 --passes 4 --precision fp16 > gpurun_out/r02d_c1_fp16_ksplit_$ks.jsonl 2>&1; echo rc=$?; cat gpurun_out/r02d_c1_fp16_ksplit_$ks.jsonl
done
unset GGRM_KSPLIT
for wt in 4 8; do
  timeout 300 python tools/cli_configs.py C4 --precision fp16 --writer-threads $wt > gpurun_out/r02d_c4_wt$wt.jsonl 2> gpurun_out/r02d_c4_wt$wt.err; echo rc=$?; cat gpurun_out/r02d_c4_wt$wt.jsonl
done
timeout 200 python tools/cli_configs.py C2 --precision bf16 > gpurun_out/r02d_c2.jsonl 2> gpurun_out/r02d_c2.err; echo rc=$?; cat gpurun_out/r02d_c2.jsonl
export GGRM_BENCH_M=131072
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02d_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-wall > gpurun_out/r02d_ncu1.log 2>&1; echo rc=$?
