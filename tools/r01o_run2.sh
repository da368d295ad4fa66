set -x
for R in pipe peer; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --reduce $R \
    > gpurun_out/r01o_bench_n2_$R.json 2> gpurun_out/r01o_bench_n2_$R.err; echo rc=$?; tail -3 gpurun_out/r01o_bench_n2_$R.err
done
python - <<'P'
import json
for r in ("pipe","peer"):
    try:
        d=json.loads(open(f"gpurun_out/r01o_bench_n2_{r}.json").read().strip().splitlines()[-1])
        print(r, "value", round(d["value"],1), "ms", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"],1), "e2e_ms", round(d["e2e"]["wall_s_per_grm"]*1e3,1), "syrk", round(d["phases_ms_per_step"]["syrk_ms"],1), "clk", d["clocks"]["sm_mhz"])
    except Exception as e: print(r, "ERR", e)
P
