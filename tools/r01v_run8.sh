set -x
python -m pytest tests/test_multiproc_gpu.py -m gpu -x -q 2>&1 | tail -2
for cfg in "8 pipe" "4 pipe" "8 peer"; do
  set -- $cfg
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29540 + $1)) bench.py --gpus $1 --steps 3 --warmup 3 --reduce $2 \
    > gpurun_out/r01v_bench_n$1_$2.json 2> gpurun_out/r01v_bench_n$1_$2.err; echo rc=$?; tail -2 gpurun_out/r01v_bench_n$1_$2.err
done
python - <<'P'
import json
for r in ("n8_pipe","n4_pipe","n8_peer"):
    try:
        d=json.loads(open(f"gpurun_out/r01v_bench_{r}.json").read().strip().splitlines()[-1])
        print(r, "value", round(d["value"],1), "ms", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"],1), "e2e_ms", round(d["e2e"]["wall_s_per_grm"]*1e3,1), "syrk", round(d["phases_ms_per_step"]["syrk_ms"],1), "clk", d["clocks"]["sm_mhz"])
    except Exception as e: print(r, "ERR", e)
P
