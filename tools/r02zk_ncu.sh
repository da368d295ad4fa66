set -x
mkdir -p gpurun_out
timeout 200 ncu --set full --clock-control none --import-source on -k regex:prep_fused -s 1 -c 1 -o gpurun_out/r02zk_fused python tools/config_probe.py C1 --precision fp16 --passes 2 > gpurun_out/r02zk_ncu1.log 2>&1; echo rc=$?
timeout 200 ncu --set full --clock-control none -k regex:gcta_diag_kernel -s 1 -c 1 -o gpurun_out/r02zk_gcta python tools/probes/gcta_diag_probe.py 20000 30000 > gpurun_out/r02zk_ncu2.log 2>&1; echo rc=$?
ls -la gpurun_out | grep r02zk
