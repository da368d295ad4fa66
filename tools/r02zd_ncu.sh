set -x
mkdir -p gpurun_out
GGRM_FUSED_PREP=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:prep_fused -s 1 -c 1 -o gpurun_out/r02zd_fused python tools/config_probe.py C1 --precision fp16 --passes 2 > gpurun_out/r02zd_ncu1.log 2>&1; echo rc=$?
GGRM_FUSED_PREP=0 timeout 300 ncu --set full --clock-control none -k regex:"expand16|repack_kernel|count_stats" -s 3 -c 3 -o gpurun_out/r02zd_unfused python tools/config_probe.py C1 --precision fp16 --passes 2 > gpurun_out/r02zd_ncu2.log 2>&1; echo rc=$?
ls -la gpurun_out | grep r02zd
