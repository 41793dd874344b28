# round 2, what the driver runs at round end, on one GPU: smoke(), both bench arms (the GPU suite: tools/gpu_r02_na.sh)
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02_bench_g1_reference.json 2> gpurun_out/r02_ref.err; echo ref rc=$?
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02_bench_g1.json 2> gpurun_out/r02_bench_g1.err; echo bench rc=$?; tail -4 gpurun_out/r02_bench_g1.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r02_bench_g1.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/r02_bench_g1_reference.json').read().strip().splitlines()[-1])
print('value', l['value'], 'e2e', l['e2e']['value'], l['e2e']['h2d_gbs_per_gpu'], 'pinned', l['e2e_pinned']['value'], 'ref', r['value'], 'same config', l['config']==r['config'], l['steps']==r['steps'], l['warmup']==r['warmup'])
print('parity', l['parity']['ok'], 'roofline', l['roofline']['frac'], l['roofline']['traffic'], 'launches', l['gpu_launches'], 'clocks', l['clocks'])
print('configs', {k: round(l[k]['value'],1) for k in l if k.startswith('configs')}, 'cpu rows', [(x['name'], round(x['value'],1)) for x in l['cpu_baseline'].get('rows',[])])
PY
