# programmatic dependent launch on/off at latency-bound sizes (one GPU)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for n in 1024 2048 4096 8192 16384; do
for d in 0 1; do
CGB_PDL=$d python bench.py --steps 400 --warmup 3 --no-e2e --no-cpu --matrix-n $n | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($n, 'pdl=$d classic', round(l['value'],1), round(l['graph_replay']['value'],1), 'sym', round(l.get('symmetric_variant',{}).get('graph_replay',0),1), 'fused', round(l.get('fused_reductions_variant',{}).get('value',0),1), 'fused+sym', round(l.get('fused_reductions_variant',{}).get('fused_symmetric',0),1))"
done
done
