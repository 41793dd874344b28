# usage: bash tools/gpu_run8.sh G   -- multi-GPU parity tests + nccl-vs-peer bench at n = 65536 and 131072
G=${1:-8}
set -x
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29611"
run() { tag=$1; shift; timeout 900 $TR bench.py --gpus $G --warmup 3 --no-cpu "$@" > gpurun_out/b${G}_$tag.json 2> gpurun_out/b${G}_$tag.err; echo "rc=$?"; python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/b${G}_$tag.json").read().strip().splitlines()[-1])
    print("G=$G $tag", "it/s", round(l["value"],1), "graph", round(l["graph_replay"]["value"],1), "gemv_ms", round(l["roofline"]["gemv_ms_avg"],4), "GB/s/GPU", round(l["roofline"]["achieved"]), "e2e", l.get("e2e",{}).get("value"), l["config"]["exchange"][:5], l["clocks"])
except Exception as e:
    print("G=$G $tag FAILED", e); print(open("gpurun_out/b${G}_$tag.err").read()[-1500:])
PY
}
run peer_65536 --matrix-n 65536 --comm peer --steps 100
run nccl_65536 --matrix-n 65536 --comm nccl --steps 100 --no-e2e
run peer_131072 --matrix-n 131072 --comm peer --steps 50 --no-e2e
run nccl_131072 --matrix-n 131072 --comm nccl --steps 50 --no-e2e
run peer_16384 --matrix-n 16384 --comm peer --steps 200 --no-e2e
run nccl_16384 --matrix-n 16384 --comm nccl --steps 200 --no-e2e
