set -x
mkdir -p gpurun_out
GGRM_DEBUG_GAPS=1 timeout 200 python tools/config_probe.py C1 --precision fp16 --passes 3 > gpurun_out/r02zg_c1.jsonl 2> gpurun_out/r02zg_c1_timeline.txt; echo rc=$?
cat gpurun_out/r02zg_c1.jsonl | cut -c1-300; grep "ggrm" gpurun_out/r02zg_c1_timeline.txt | tail -24
