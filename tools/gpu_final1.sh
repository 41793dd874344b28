# round-end check on one GPU: full GPU test suite, smoke, default bench (both arms)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/final_default.json 2> gpurun_out/final_default.err; echo rc=$?
tail -1 gpurun_out/final_default.json | python -c "import json,sys; l=json.loads(sys.stdin.read()); print({k: l[k] for k in ('value','ms_per_step','e2e','roofline','clocks','gpu_launches')}); print(l.get('symmetric_variant')); print(l.get('fused_reductions_variant')); print(l.get('cpu_baseline'))"
timeout 900 python bench.py --impl reference > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err; echo rc=$?
tail -1 gpurun_out/final_reference.json
