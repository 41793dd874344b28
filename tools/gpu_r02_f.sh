# usage: bash tools/gpu_r02_f.sh G  -- probe of the exchange kernels, H2D modes with G concurrent uploads, new multi-GPU tests
G=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29631"
timeout 200 python -m pytest tests/test_gpu_multi.py -x -q -k "merged or pinned_in_place" 2>&1 | tail -4
for n in 65536 16384; do
  timeout 120 $TR tools/probe_exchange.py --order $n > gpurun_out/r02_probe_g${G}_n$n.json 2> gpurun_out/r02_probe_g${G}_n$n.err; echo probe rc=$?
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02_probe_g${G}_n$n.json').read().strip().splitlines()[-1])
for r in d['ranks'][:2]+d['ranks'][-1:]:
    print(json.dumps(r)[:1400])
PY
done
timeout 300 $TR tools/h2d_ceiling.py --order 16384 --reps 3 > gpurun_out/r02_h2d_g$G.json 2> gpurun_out/r02_h2d_g$G.err; echo h2d rc=$?
cut -c1-2500 gpurun_out/r02_h2d_g$G.json; tail -2 gpurun_out/r02_h2d_g$G.err | cut -c1-300
