# Developer probe: DRAM bytes per SYRK launch for item-order / slab-length variants (ncu, n=50000).
run() { echo "== $*"; env "$@" ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:syrk_tc -s 2 -c 2 python tools/perf_probe.py 50000 ${M:-32768} 2 0 1 2>&1 | grep -E "dram__|gpu__time|hit_rate" ; }
run GGRM_STRIP_H=0
run GGRM_STRIP_H=4
run GGRM_STRIP_H=6
run GGRM_STRIP_H=8
run GGRM_STRIP_H=9
run GGRM_STRIP_H=12
run GGRM_STRIP_H=16
run GGRM_STRIP_H=8 GGRM_CACHE_HINTS=0
run GGRM_STRIP_H=8 GGRM_CACHE_HINTS=1
run GGRM_STRIP_H=8 GGRM_CACHE_HINTS=2
M=65536 run GGRM_STRIP_H=8 GGRM_CHUNK_SNPS=16384
M=16384 run GGRM_STRIP_H=8 GGRM_CHUNK_SNPS=4096
