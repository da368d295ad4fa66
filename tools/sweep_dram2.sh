run() { echo "== $*"; env "$@" ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum --clock-control none -k regex:syrk_tc -s 2 -c 2 python tools/perf_probe.py 50000 ${M:-32768} 2 0 1 2>&1 | grep -E "dram__|gpu__time" ; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
GGRM_WAVE_SYNC=1 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
run GGRM_STRIP_H=8 GGRM_WAVE_SYNC=1
run GGRM_STRIP_H=8 GGRM_WAVE_SYNC=2
run GGRM_STRIP_H=8 GGRM_WAVE_SYNC=4
run GGRM_STRIP_H=12 GGRM_WAVE_SYNC=1
run GGRM_STRIP_H=0 GGRM_WAVE_SYNC=1
M=65536 run GGRM_STRIP_H=8 GGRM_CHUNK_SNPS=16384 GGRM_WAVE_SYNC=1
t() { echo "== T $*"; env "$@" python tools/perf_probe.py 50000 196608 2 0 4 2>&1 | grep pass; }
t GGRM_STRIP_H=8 GGRM_WAVE_SYNC=0
t GGRM_STRIP_H=8 GGRM_WAVE_SYNC=1
t GGRM_STRIP_H=8 GGRM_WAVE_SYNC=2
t GGRM_STRIP_H=12 GGRM_WAVE_SYNC=1
t GGRM_STRIP_H=8 GGRM_WAVE_SYNC=1 GGRM_CHUNK_SNPS=16384
t GGRM_STRIP_H=8 GGRM_WAVE_SYNC=0
