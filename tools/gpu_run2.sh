set -x
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
for n in 4096 16384 65536; do
timeout 600 $TR bench.py --gpus 2 --steps 200 --warmup 3 --no-e2e --no-cpu --matrix-n $n > gpurun_out/b2f_$n.json 2> gpurun_out/b2f_$n.err; echo rc=$?
python -c "import json; l=json.loads(open('gpurun_out/b2f_$n.json').read().strip().splitlines()[-1]); print($n, 'classic', round(l['value'],1), round(l['graph_replay']['value'],1), 'sym', l.get('symmetric_variant',{}).get('graph_replay'), 'fused', l.get('fused_reductions_variant',{}).get('value'), l.get('fused_reductions_variant',{}).get('fused_symmetric'))" || tail -20 gpurun_out/b2f_$n.err
done
