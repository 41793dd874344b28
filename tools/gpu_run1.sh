set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
run() { tag=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/b1_$tag.json 2> gpurun_out/b1_$tag.err; echo "rc=$?"; python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/b1_$tag.json").read().strip().splitlines()[-1])
    print("G=1 $tag", "it/s", round(l["value"],2), "graph", round(l["graph_replay"]["value"],2), "gemv_ms", round(l["roofline"]["gemv_ms_avg"],4), "GB/s", round(l["roofline"]["achieved"]), "e2e", l.get("e2e",{}).get("value"), "cpu", l.get("cpu_baseline",{}).get("value"), l["clocks"])
except Exception as e:
    print("G=1 $tag FAILED", e); print(open("gpurun_out/b1_$tag.err").read()[-1500:])
PY
}
run default
run n131072 --matrix-n 131072 --steps 20 --no-e2e --no-cpu
run n16384 --matrix-n 16384 --steps 200 --no-e2e --no-cpu
run n4096 --matrix-n 4096 --steps 2000 --no-e2e --no-cpu
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/b1_reference.json 2>gpurun_out/b1_reference.err; tail -c 600 gpurun_out/b1_reference.json
