# usage: bash tools/gpu_r02_g.sh G -- staging chunk sweep (real small chunks this time) and, G > 1, the early-flag product + forced PDL
G=${1:-1}
mkdir -p gpurun_out
if [ "$G" = 1 ]; then
  timeout 600 python tools/h2d_ceiling.py --order 24576 --sweep > gpurun_out/r02_h2d_g1_sweep.json 2> gpurun_out/r02_h2d_g1_sweep.err; echo rc=$?
  cut -c1-3000 gpurun_out/r02_h2d_g1_sweep.json
  timeout 300 python -m pytest tests/test_gpu_staging.py -x -q 2>&1 | tail -3
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29641"
  timeout 200 python -m pytest tests/test_gpu_multi.py -x -q -k "generated_sharded or long_running or merged or staging_ring" 2>&1 | tail -3
  run() {  # n pdl
    CGB_PDL=$2 timeout 200 $TR bench.py --gpus $G --warmup 3 --steps 200 --no-e2e --no-cpu --no-parity --no-configs --no-variants --matrix-n $1 2>gpurun_out/ab.err | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.read())
print('G=$G n=$1 pdl=$2', 'value', round(l['value'],1), 'us/it', round(l['ms_per_step']*1e3,2), 'instr', round(l['instrumented_pass']['value'],1), 'gemv_us', round(l['roofline']['gemv_ms_avg']*1e3,2), 'GB/s', round(l['roofline']['achieved'],0), 'pdl', l['implementation']['launch'])" || tail -3 gpurun_out/ab.err
  }
  for n in 65536 16384; do run $n 1; run $n 2; run $n 1; run $n 2; done
  timeout 120 $TR tools/probe_exchange.py --order 65536 > gpurun_out/r02_probe2_g${G}_n65536.json 2> gpurun_out/r02_probe2.err; echo probe rc=$?
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02_probe2_g${G}_n65536.json').read().strip().splitlines()[-1])
for r in d['ranks'][:1]: print(json.dumps(r)[:1400])
PY
  timeout 300 $TR tools/h2d_ceiling.py --order 16384 --reps 3 > gpurun_out/r02_h2d_g$G.json 2> gpurun_out/r02_h2d_g$G.err; echo h2d rc=$?
  cut -c1-2500 gpurun_out/r02_h2d_g$G.json
fi
