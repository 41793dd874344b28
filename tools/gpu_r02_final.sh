# usage: bash tools/gpu_r02_final.sh G [tests]   -- round 2 closing run on G GPUs: multi-GPU parity tests, the full bench line, the H2D ceiling, the exchange probes
G=${1:-8}
mkdir -p gpurun_out
nproc; free -g | head -2 | tail -1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29611"
if [ "${2:-tests}" = tests ]; then
  ( time timeout 600 python -m pytest tests/test_gpu_multi.py -x -q ) > gpurun_out/r02_multi_test_g$G.log 2>&1; tail -4 gpurun_out/r02_multi_test_g$G.log
fi
( time timeout 400 $TR bench.py --gpus $G --steps 20 --warmup 5 ) > gpurun_out/r02_bench_g$G.json 2> gpurun_out/r02_bench_g$G.err; echo bench rc=$?
python - <<PY
import json
l=json.loads(open('gpurun_out/r02_bench_g$G.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','parity','e2e','e2e_pinned','configs4_n131072','configs2_n16384','roofline','symmetric_variant','fused_reductions_variant'):
    print(k, json.dumps(l.get(k))[:700])
PY
timeout 200 $TR tools/h2d_ceiling.py --order 16384 --reps 3 > gpurun_out/r02_h2d_g$G.json 2> gpurun_out/r02_h2d_g$G.err; echo h2d rc=$?
cut -c1-2200 gpurun_out/r02_h2d_g$G.json
for n in 65536 16384; do
  timeout 120 $TR tools/probe_exchange.py --order $n > gpurun_out/r02_probe_g${G}_n$n.json 2> gpurun_out/r02_probe_g${G}_n$n.err; echo probe $n rc=$?
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02_probe_g${G}_n$n.json').read().strip().splitlines()[-1])
for r in d['ranks'][:1]: print(json.dumps(r)[:1500])
PY
done
