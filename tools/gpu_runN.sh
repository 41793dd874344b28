# usage: bash tools/gpu_runN.sh G "n1 n2 ..."  -- multi-GPU parity tests, then bench (default exchange, symmetric variant included)
G=$1; SIZES=$2
set -x
timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_ingest.py tests/test_gpu_guards.py -m gpu -x -q 2>&1 | tail -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29611"
for n in $SIZES; do
  steps=100; [ $n -ge 131072 ] && steps=50
  E2E="--no-e2e"; [ $n -eq 65536 ] && E2E=""        # the headline size runs the driver's default legs (e2e included)
  timeout 900 $TR bench.py --gpus $G --warmup 3 --steps $steps --no-cpu $E2E --matrix-n $n > gpurun_out/bs${G}_$n.json 2> gpurun_out/bs${G}_$n.err; echo rc=$?
  python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/bs${G}_$n.json").read().strip().splitlines()[-1]); s=l.get("symmetric_variant") or {}
    print("G=$G n=$n it/s", round(l["value"],1), "graph", round(l["graph_replay"]["value"],1), "gemv_ms", round(l["roofline"]["gemv_ms_avg"],4), "| symmetric it/s", round(s.get("value",0),1), "graph", round(s.get("graph_replay",0),1), "gemv_ms", round(s.get("gemv_ms_avg",0),4), "GB/s declared", round(s.get("achieved_gbs_on_declared_bytes",0)), l["clocks"], "e2e", (l.get("e2e") or {}).get("value"), "sym e2e", ((s.get("e2e") or {}).get("value")), "fused", (l.get("fused_reductions_variant") or {}).get("value"), (l.get("fused_reductions_variant") or {}).get("fused_symmetric"))
except Exception as e:
    print("FAILED", e); print(open("gpurun_out/bs${G}_$n.err").read()[-1500:])
PY
done
