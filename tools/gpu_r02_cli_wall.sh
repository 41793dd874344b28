# where the CLI's wall time goes on a box whose only GPU client it is: bare CUDA start-up against the whole program
mkdir -p gpurun_out
nvidia-smi --query-gpu=persistence_mode --format=csv,noheader
for i in 1 2 3; do tools/bin/cuda_init_time; done
P=/tmp/sysw
conjugate_gradient_b200/lib/cg_b200 --gen 8192 --seed 20170209 --emit-files $P > /dev/null
TIMEFORMAT="   wall %R s user %U s sys %S s"
for i in 1 2 3; do
  echo "cg_b200 on files:"; time ( CGB_INGEST_TIMING=1 conjugate_gradient_b200/lib/cg_b200 ${P}_matrix.txt ${P}_vector.txt ${P}_x0.txt --dims ${P}_dimensions.txt --stats 2>&1 | tail -12 )
done
echo "cg_b200 --gen 8192 (no files):"; time ( conjugate_gradient_b200/lib/cg_b200 --gen 8192 --seed 20170209 2>&1 | tail -3 )
