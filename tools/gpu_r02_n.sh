# round 2, 1 GPU: ncu evidence -- launch list and full capture of the headline product, the per-GPU slice shapes, the R R' generator
mkdir -p gpurun_out
B="python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-parity --no-configs --no-variants"
$B > gpurun_out/r02_ncu_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_n65536.csv $B > gpurun_out/r02_ncu_l.log 2>&1; echo launches rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cgb_rowblock_matvec_ldg -s 6 -c 2 -f -o gpurun_out/r02_gemv_n65536 $B > gpurun_out/r02_ncu_f.log 2>&1; echo full rc=$?
for m in 8192 16384; do
  S="python tools/gemv_slice.py $m 65536 2 4 2 256"
  $S > gpurun_out/r02_slice_$m.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:cgb_rowblock_matvec -c 2 -f -o gpurun_out/r02_slice_${m}x65536 $S > gpurun_out/r02_ncu_s$m.log 2>&1; echo slice $m rc=$?; cat gpurun_out/r02_slice_$m.log
done
G="python -c \"import conjugate_gradient_b200 as cg; c = cg.Context(0); c.generate(8192, 1, cg.GEN_GRAM); c.close()\""
eval $G && eval timeout 600 ncu --set full --clock-control none --import-source on -k regex:gram -c 1 -f -o gpurun_out/r02_gram_n8192 $G > gpurun_out/r02_ncu_g.log 2>&1; echo gram rc=$?
for f in r02_gemv_n65536 r02_slice_8192x65536 r02_slice_16384x65536 r02_gram_n8192; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_ncu_raw.csv 2>/dev/null
  python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/${f}_ncu_raw.csv'))); h=rows[0]
for k in ('Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__throughput.avg.pct_of_peak_sustained_elapsed','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fp64.sum','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active'):
    if k in h: print('$f', k, [r[h.index(k)][:60] for r in rows[2:]])
PY
done
ls -la gpurun_out/*.ncu-rep | awk '{print $5, $9}'
