mkdir -p gpurun_out
timeout 600 python tools/sweep_rows.py > gpurun_out/r02_rows_sweep.txt 2>&1; cat gpurun_out/r02_rows_sweep.txt
