# Third GPU call of round 2 (1 GPU): operand-split thresholds, K rotation + longer slabs for few-tile problems, CLI fast exit,
# LOCO output preparer; the whole GPU suite; both bench configs; CLI walls of C1/C2/C3/C4 with the start-up breakdown; round-2 profiles.
#   gpurun --timeout 2400 -- 'bash tools/r02c_run.sh'
set -x
mkdir -p gpurun_out
df -h /dev/shm /tmp > gpurun_out/r02c_df.txt 2>&1; nproc >> gpurun_out/r02c_df.txt; free -g >> gpurun_out/r02c_df.txt
timeout 900 python -m pytest tests -m gpu -q -x -s > gpurun_out/r02c_pytest.log 2>&1; echo rc=$?; grep -E "max abs|max \||unsplit|passed|failed|Error" gpurun_out/r02c_pytest.log | tail -40
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r02c_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r02c_smoke.log
for ks in 1 0; do
  if [ $ks = 1 ]; then export GGRM_KSPLIT=1; else unset GGRM_KSPLIT; fi
  timeout 120 python tools/config_probe.py C1 --passes 4 > gpurun_out/r02c_c1_ksplit_$ks.jsonl 2>&1; echo rc=$?; cat gpurun_out/r02c_c1_ksplit_$ks.jsonl
done
unset GGRM_KSPLIT
timeout 400 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r02c_bench_c2.json 2> gpurun_out/r02c_bench_c2.err; echo rc=$?; cat gpurun_out/r02c_bench_c2.json; tail -5 gpurun_out/r02c_bench_c2.err
GGRM_DEBUG_TIMING=1 timeout 300 python tools/cli_configs.py C1 C3 C2 --precision fp16 > gpurun_out/r02c_cli.jsonl 2> gpurun_out/r02c_cli.err; echo rc=$?; cat gpurun_out/r02c_cli.jsonl
timeout 600 python bench.py --config c4 --wall --gpus 1 --steps 2 --warmup 2 --no-cpu > gpurun_out/r02c_bench_c4.json 2> gpurun_out/r02c_bench_c4.err; echo rc=$?; cat gpurun_out/r02c_bench_c4.json; tail -5 gpurun_out/r02c_bench_c4.err
export GGRM_BENCH_M=131072
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02c_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-wall > gpurun_out/r02c_ncu1.log 2>&1; echo rc=$?
timeout 200 ncu --set full --clock-control none --import-source on -k regex:pack_rows8 -s 2 -c 1 -o gpurun_out/r02c_pack python bench.py --steps 1 --warmup 1 --no-cpu --no-wall > gpurun_out/r02c_ncu2.log 2>&1; echo rc=$?
ls -la gpurun_out | grep r02c
