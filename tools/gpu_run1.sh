set -x
CMD="python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/ncu_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01b_launches_n65536.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
$CMD > gpurun_out/ncu_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemv_ldg_kernel -s 2 -c 2 -o gpurun_out/r01b_gemv_n65536 $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
CMD4="python bench.py --steps 200 --warmup 3 --no-e2e --no-cpu --matrix-n 4096"
$CMD4 > gpurun_out/ncu_plain4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemv_ldg_keep -s 20 -c 2 -o gpurun_out/r01b_gemv_keep_n4096 $CMD4 > gpurun_out/ncu_full4.log 2>&1
echo "full4 rc=$?"
tail -c 400 gpurun_out/ncu_plain4.log
ls -la gpurun_out/*.ncu-rep
python tools/sweep_l2.py 8192
