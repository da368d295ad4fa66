# First GPU call of the next round: everything written after round 1's GPU minutes were spent, before anything else is built on it.
#   gpurun --timeout 1500 -- 'bash tools/r02_first_run.sh'
# 1. the regular GPU suite + smoke (writer Auto mode is on by default: the engine tests now write through the shared mapping
#    for tile rows), 2. the opt-in BED relay test, 3. bench + reference arm, 4. CLI walls of C2 / C3 with both writers.
set -x
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; echo rc=$?; tail -5 gpurun_out/r02a_pytest.log
GELEX_B200_TEST_UNVERIFIED=1 timeout 300 python -m pytest tests -m gpu -q -k "bed_relay or writer_modes or random_small_inputs or reference_pipe_files or reappearing or reference_host_code or prefault or stage_slots" > gpurun_out/r02a_unverified.log 2>&1; echo rc=$?; tail -8 gpurun_out/r02a_unverified.log
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r02a_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r02a_smoke.log
timeout 300 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo rc=$?; cat gpurun_out/r02a_bench.json
timeout 300 python bench.py --impl reference --gpus 1 --steps 3 --warmup 3 > gpurun_out/r02a_reference.json 2> gpurun_out/r02a_reference.err; echo rc=$?; cat gpurun_out/r02a_reference.json
for w in pwrite mmap; do
  GELEX_B200_WRITER=$w timeout 300 python tools/cli_configs.py C3 C2 > gpurun_out/r02a_cli_$w.jsonl 2> gpurun_out/r02a_cli_$w.err; echo rc=$?; cat gpurun_out/r02a_cli_$w.jsonl
done
GELEX_B200_WRITER=mmap GELEX_B200_PREFAULT=1 timeout 300 python tools/cli_configs.py C3 C2 > gpurun_out/r02a_cli_prefault.jsonl 2> gpurun_out/r02a_cli_prefault.err; echo rc=$?; cat gpurun_out/r02a_cli_prefault.jsonl
python tools/writer_probe.py --n 36000 --dirs /dev/shm /tmp > gpurun_out/r02a_writer_probe.jsonl 2>&1; tail -24 gpurun_out/r02a_writer_probe.jsonl
