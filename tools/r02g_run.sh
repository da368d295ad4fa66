# Seventh GPU call (1 GPU): short first slab, final strip as its own band, lazy writer pools, smaller pinned ring - whole GPU suite,
# bench C2 at N = 1, C1 probe, C2 CLI.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02g_pytest.log 2>&1; echo rc=$?; tail -6 gpurun_out/r02g_pytest.log
timeout 120 python tools/config_probe.py C1 C3 --passes 4 --precision fp16 > gpurun_out/r02g_c1c3.jsonl 2>&1; echo rc=$?; cat gpurun_out/r02g_c1c3.jsonl
timeout 400 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu > gpurun_out/r02g_bench_c2.json 2> gpurun_out/r02g_bench_c2.err; echo rc=$?; cat gpurun_out/r02g_bench_c2.json; tail -5 gpurun_out/r02g_bench_c2.err
