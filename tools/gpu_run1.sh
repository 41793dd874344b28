timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused" 2>&1 | tail -3
for n in 2048 4096 8192; do
python bench.py --steps 400 --warmup 3 --no-e2e --no-cpu --matrix-n $n | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($n, 'classic', round(l['value'],1), round(l['graph_replay']['value'],1), 'sym', l.get('symmetric_variant',{}).get('graph_replay'), 'fused', l.get('fused_reductions_variant',{}).get('value'), l.get('fused_reductions_variant',{}).get('fused_symmetric'))"
done
