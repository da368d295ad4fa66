set -x
mkdir -p gpurun_out
timeout 400 python tools/probes/gcta_diag_probe.py 50000 500000 > gpurun_out/r02ze_gcta_c2.jsonl 2> gpurun_out/r02ze.err; echo rc=$?
cat gpurun_out/r02ze_gcta_c2.jsonl; tail -5 gpurun_out/r02ze.err
