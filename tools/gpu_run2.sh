set -x
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "symmetric" 2>&1 | tail -15
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
for n in 16384 65536; do
timeout 600 $TR bench.py --gpus 2 --steps 40 --warmup 3 --no-e2e --no-cpu --matrix-n $n > gpurun_out/b2s_$n.json 2> gpurun_out/b2s_$n.err; echo rc=$?
python -c "import json; l=json.loads(open('gpurun_out/b2s_$n.json').read().strip().splitlines()[-1]); print($n, l['value'], l['graph_replay']['value'], l.get('symmetric_variant'))" || tail -20 gpurun_out/b2s_$n.err
done
