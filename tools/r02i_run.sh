# C1 after the few-tile slab rule (one slab, K parts so that a TMEM chain stays <= 16 384 SNPs) and the flattened repack; whole GPU
# suite (repack and split-K are shared code); bench C2 N = 1 as the regression check.
set -x
mkdir -p gpurun_out
GGRM_DEBUG_GAPS=1 timeout 120 python tools/config_probe.py C1 C3 --passes 6 --precision fp16 > gpurun_out/r02i_c1c3.jsonl 2> gpurun_out/r02i_c1c3.err; echo rc=$?; cat gpurun_out/r02i_c1c3.jsonl
timeout 120 python tools/config_probe.py C1 --passes 6 --precision bf16 > gpurun_out/r02i_c1_bf16.jsonl 2>&1; echo rc=$?; cat gpurun_out/r02i_c1_bf16.jsonl
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02i_pytest.log 2>&1; echo rc=$?; tail -5 gpurun_out/r02i_pytest.log
timeout 400 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu > gpurun_out/r02i_bench_c2.json 2> gpurun_out/r02i_bench_c2.err; echo rc=$?; cut -c1-400 gpurun_out/r02i_bench_c2.json
