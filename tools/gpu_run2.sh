set -x
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
run() { tag=$1; shift; timeout 600 $TR bench.py --gpus 2 --steps 40 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/b2_$tag.json 2> gpurun_out/b2_$tag.err; echo "rc=$?"; python - <<PY
import json
l=json.loads(open("gpurun_out/b2_$tag.json").read().strip().splitlines()[-1])
print("$tag", "it/s", round(l["value"],1), "graph", round(l["graph_replay"]["value"],1), "gemv_ms", round(l["roofline"]["gemv_ms_avg"],4), "GB/s", round(l["roofline"]["achieved"]), l["config"]["gemv_kernel"]["gemv_rows"], l["config"]["gemv_kernel"]["gemv_threads"], l["config"]["exchange"][:5])
PY
}
run nccl_16384 --matrix-n 16384 --comm nccl
run peer_16384 --matrix-n 16384 --comm peer
run peer_16384_r4 --matrix-n 16384 --comm peer --rows 4 --unroll 2 --threads 256
run nccl_65536 --matrix-n 65536 --comm nccl
run nccl_65536_r4 --matrix-n 65536 --comm nccl --rows 4 --unroll 2 --threads 256
run peer_65536 --matrix-n 65536 --comm peer
run peer_65536_r2 --matrix-n 65536 --comm peer --rows 2 --unroll 2 --threads 256
run peer_4096 --matrix-n 4096 --comm peer
run nccl_4096 --matrix-n 4096 --comm nccl
