# usage: bash tools/gpu_r02_c.sh G [tests]   -- round 2: merged vector kernel and (8,2,512) exchange-fused product, A/B on G GPUs
G=${1:-2}
mkdir -p gpurun_out
if [ "${2:-tests}" = tests ]; then
  ( time timeout 1500 python -m pytest tests/test_gpu_multi.py -x -q ) 2>&1 | tail -6
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29611"
run() {  # n xrp rows unroll threads
  CGB_XRP=$2 timeout 600 $TR bench.py --gpus $G --warmup 3 --steps 200 --no-e2e --no-cpu --no-parity --no-configs --no-variants --matrix-n $1 --rows $3 --unroll $4 --threads $5 2>gpurun_out/ab.err | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.read())
print('G=$G n=$1 xrp=$2 shape=($3,$4,$5)', 'value', round(l['value'],1), 'us/it', round(l['ms_per_step']*1e3,2), 'instr', round(l['instrumented_pass']['value'],1), 'gemv_us', round(l['roofline']['gemv_ms_avg']*1e3,2), 'GB/s', round(l['roofline']['achieved'],0), 'k/it', l['timed_region']['kernels_per_iteration'], l['implementation']['gemv_kernel']['gemv_rows'], l['implementation']['gemv_kernel']['gemv_threads'])" || tail -3 gpurun_out/ab.err
}
for n in 65536 16384; do
  run $n 0 0 0 0
  run $n 1 0 0 0
  run $n 0 8 2 512
  run $n 1 8 2 512
  run $n 1 0 0 0
  run $n 1 8 2 512
done
