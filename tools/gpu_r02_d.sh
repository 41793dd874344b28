# usage: bash tools/gpu_r02_d.sh G   -- which ncu metric sets are single-pass on the exchange-fused kernels (rank 0 under ncu)
G=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29621 --no-python"
ARGS="--gpus $G --warmup 3 --steps 20 --no-e2e --no-cpu --no-parity --no-configs --no-variants --matrix-n 65536"
i=0
for M in "gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum" "nvlrx__bytes.sum,nvltx__bytes.sum" "gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,nvlrx__bytes.sum,nvltx__bytes.sum,lts__t_sectors_srcunit_tex_op_read.sum"; do
  i=$((i+1))
  timeout 300 $TR tools/ncu_rank0.sh gpurun_out/r02_ncu_g${G}_set$i "$M" $ARGS > gpurun_out/r02_ncu_g${G}_set$i.log 2>&1; echo "set$i rc=$?"
  grep -c "cgb_" gpurun_out/r02_ncu_g${G}_set$i.csv; grep -m3 "passes\|==ERROR\|==WARNING" gpurun_out/r02_ncu_g${G}_set$i.csv gpurun_out/r02_ncu_g${G}_set$i.log | head -5
  tail -2 gpurun_out/r02_ncu_g${G}_set$i.log | cut -c1-300
done
ncu --query-metrics 2>/dev/null | grep -i "nvl" | head -20
