# comm-GEMV shape sweep on 2 GPUs: usage gpu_sweep_comm.sh "n1 n2"
for n in $1; do
for cfg in "4 2 256" "2 1 256" "2 2 256" "4 1 128"; do
set -- $cfg
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 100 --warmup 3 --no-e2e --no-cpu --matrix-n $n --rows $1 --unroll $2 --threads $3 2>/dev/null | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($n, 'cfg', '$cfg', round(l['value'],1), round(l['graph_replay']['value'],1), 'gemv_ms', round(l['roofline']['gemv_ms_avg'],4), l['config']['gemv_kernel']['gemv_rows'], l['config']['gemv_kernel']['gemv_unroll'], l['config']['gemv_kernel']['gemv_threads'])"
done
done
