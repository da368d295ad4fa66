# C1 (2 000 x 50 000, FP16, BED resident): slab cuts for few-tile problems.  default = 2 slabs of 25 024 (split-K 2);
# then a short first slab in front, one single slab, four slabs; GGRM_DEBUG_GAPS prints idle time between SYRK launches.
set -x
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" GGRM_DEBUG_GAPS=1 timeout 120 python tools/config_probe.py C1 --passes 6 --precision fp16 > gpurun_out/r02h_c1_$tag.jsonl 2> gpurun_out/r02h_c1_$tag.err; echo rc=$?; cat gpurun_out/r02h_c1_$tag.jsonl; grep gaps gpurun_out/r02h_c1_$tag.err | tail -2; }
run default X=1
run first4096 GGRM_FIRST_SLAB_SPLIT=1
run first2048 GGRM_FIRST_SLAB_SPLIT=1 GGRM_FIRST_SLAB=2048
run first8192 GGRM_FIRST_SLAB_SPLIT=1 GGRM_FIRST_SLAB=8192
run one_slab GGRM_CHUNK_SNPS=50048
run four_slabs GGRM_CHUNK_SNPS=12544
run ksplit1_one_slab GGRM_KSPLIT=1 GGRM_CHUNK_SNPS=50048
run serial_prep GGRM_OVERLAP_PREP=0
