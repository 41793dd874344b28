# round 2, closing run on one GPU: the GPU suite, smoke(), both bench arms as the driver runs them, whole programs on files, ingest kernels' launch list
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gpu_suite_close.log 2>&1; grep -E "passed|failed|error" gpurun_out/r02_gpu_suite_close.log | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02_close_bench_g1_reference.json 2> gpurun_out/r02_close_ref.err; echo ref rc=$?; grep real gpurun_out/r02_close_ref.err
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02_close_bench_g1.json 2> gpurun_out/r02_close_bench_g1.err; echo bench rc=$?; grep real gpurun_out/r02_close_bench_g1.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r02_close_bench_g1.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/r02_close_bench_g1_reference.json').read().strip().splitlines()[-1])
print('value', l['value'], 'e2e', l['e2e']['value'], l['e2e']['h2d_gbs_per_gpu'], 'pinned', l['e2e_pinned']['value'], 'ref', r['value'], 'same config', l['config']==r['config'], l['steps']==r['steps'], l['warmup']==r['warmup'])
print('parity', l['parity']['ok'], 'roofline', l['roofline']['frac'], l['roofline']['traffic'], 'launches', l['gpu_launches'], 'clocks', l['clocks'])
PY
timeout 400 python tools/whole_program_table.py 8192 > gpurun_out/r02_whole_program_n8192.md 2> gpurun_out/r02_whole_program.err; echo whole rc=$?; cat gpurun_out/r02_whole_program_n8192.md
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ingest -c 90 --csv --log-file gpurun_out/r02_ingest_launches.csv python tools/ingest_rate.py 2048 8 > gpurun_out/r02_ingest_ncu.log 2>&1; echo ncu rc=$?; tail -4 gpurun_out/r02_ingest_launches.csv
