# usage: bash tools/gpu_r02_first.sh G   (G = 1, or the rank count of a multi-GPU box)
# First GPU call of the next round: confirms what was added after round 1's last GPU session.
#   1. full GPU suite (new since the last GPU session: tests/test_gpu_reference_callers.py -- C callers, stdout contract, long-run golden)
#   2. default bench, both arms
#   3. G > 1: e2e upload rate with and without the NUMA-local placement of the pinned host buffers
#      (bench.py NumaLocal; CGB_BENCH_NUMA=0 is the A/B switch) -- look at e2e.h2d_gbs_per_gpu
G=${1:-1}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
if [ "$G" -eq 1 ]; then
  timeout 900 python bench.py > gpurun_out/r02_default.json 2> gpurun_out/r02_default.err; echo rc=$?
  timeout 900 python bench.py --impl reference > gpurun_out/r02_reference.json 2> gpurun_out/r02_reference.err; echo rc=$?
  tail -1 gpurun_out/r02_default.json | python -c "import json,sys; l=json.loads(sys.stdin.read()); print({k: l[k] for k in ('value','e2e','roofline','clocks','gpu_launches')})"
  tail -1 gpurun_out/r02_reference.json
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29611"
  for numa in 1 0; do
    CGB_BENCH_NUMA=$numa timeout 900 $TR bench.py --gpus $G --warmup 3 --steps 100 --no-cpu > gpurun_out/r02_g${G}_numa$numa.json 2> gpurun_out/r02_g${G}_numa$numa.err; echo rc=$?
    tail -1 gpurun_out/r02_g${G}_numa$numa.json | python -c "import json,sys; l=json.loads(sys.stdin.read()); print('numa=$numa', l['value'], l['e2e'])"
  done
  lscpu | grep -i "numa\|socket\|model name"; nvidia-smi topo -m | head -20
fi
# 4. G = 1 only: ncu capture of the per-GPU slice shape of the 8-GPU headline run (8192 x 65536)
if [ "$G" -eq 1 ]; then
  timeout 600 python tools/gemv_slice.py && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_ldg -c 2 -f -o gpurun_out/r02_slice_8192x65536 python tools/gemv_slice.py > gpurun_out/r02_slice_ncu.log 2>&1
  ncu -i gpurun_out/r02_slice_8192x65536.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
for k in ('gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):
    if k in h: print(k, [r[h.index(k)] for r in rows[2:]])
" | tee gpurun_out/r02_slice_summary.txt
fi
