set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x -k "in_place or fused or relay or push_device or no_alloc or split_k or gapped" > gpurun_out/r02zh_tests.log 2>&1; echo rc=$?; tail -6 gpurun_out/r02zh_tests.log
for d in 0 1 0 1; do
  GGRM_DIRECT_SRC=$d GGRM_DEBUG_GAPS=1 timeout 200 python tools/config_probe.py C1 --precision fp16 --passes 4 2> gpurun_out/r02zh_timeline$d.txt | tee -a gpurun_out/r02zh_c1_direct$d.jsonl | cut -c1-260
done
grep ggrm gpurun_out/r02zh_timeline1.txt | tail -7
