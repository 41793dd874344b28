# usage: bash tools/gpu_r02_e.sh G   -- round 2 on G GPUs: multi-GPU parity tests, A/B of the merged vector kernel, the full bench line, the H2D ceiling
G=${1:-8}
mkdir -p gpurun_out
nproc; free -g | head -2 | tail -1
( time timeout 500 python -m pytest tests/test_gpu_multi.py -x -q ) > gpurun_out/r02_multi_test_g$G.log 2>&1; tail -4 gpurun_out/r02_multi_test_g$G.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29611"
run() {  # n xrp
  CGB_XRP=$2 timeout 200 $TR bench.py --gpus $G --warmup 3 --steps 200 --no-e2e --no-cpu --no-parity --no-configs --no-variants --matrix-n $1 2>gpurun_out/ab.err | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.read())
print('G=$G n=$1 xrp=$2', 'value', round(l['value'],1), 'us/it', round(l['ms_per_step']*1e3,2), 'instr', round(l['instrumented_pass']['value'],1), 'gemv_us', round(l['roofline']['gemv_ms_avg']*1e3,2), 'GB/s', round(l['roofline']['achieved'],0), 'k/it', l['timed_region']['kernels_per_iteration'])" || tail -3 gpurun_out/ab.err
}
for n in 65536 16384; do run $n 0; run $n 1; run $n 0; run $n 1; done
( time timeout 400 $TR bench.py --gpus $G --steps 20 --warmup 5 ) > gpurun_out/r02_bench_g$G.json 2> gpurun_out/r02_bench_g$G.err; echo bench rc=$?
tail -c 300 gpurun_out/r02_bench_g$G.err
python - <<PY
import json
l=json.loads(open('gpurun_out/r02_bench_g$G.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','parity','e2e','e2e_pinned','configs4_n131072','configs2_n16384','roofline'):
    print(k, json.dumps(l.get(k))[:900])
PY
timeout 300 $TR tools/h2d_ceiling.py --n 16384 --reps 3 > gpurun_out/r02_h2d_g$G.json 2> gpurun_out/r02_h2d_g$G.err; echo h2d rc=$?
cat gpurun_out/r02_h2d_g$G.json | cut -c1-1500
nvidia-smi topo -m 2>/dev/null | head -12
