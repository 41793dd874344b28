# round 2, GPU call B (1 GPU): new bench line (parity, e2e on malloc'ed buffers, configs blocks, CPU rows) + the n = 65536 oracle test
mkdir -p gpurun_out
( time timeout 1500 python bench.py --steps 20 --warmup 5 ) > gpurun_out/r02_bench_g1.json 2> gpurun_out/r02_bench_g1.err; echo bench rc=$?
tail -c 600 gpurun_out/r02_bench_g1.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r02_bench_g1.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','parity','e2e','e2e_pinned','configs4_n131072','configs2_n16384','configs1_n4096','cpu_baseline','roofline','time_to_solution'):
    print(k, json.dumps(l.get(k))[:1500])
PY
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err; echo ref rc=$?
cat gpurun_out/r02_bench_ref.json | cut -c1-900
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "baseline_size" ) 2>&1 | tail -8
