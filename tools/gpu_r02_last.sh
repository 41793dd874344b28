# round 2, last GPU call: the suite on the final library (text writer changed after the closing run), the CPU-beside-GPU table
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gpu_suite_last.log 2>&1; grep -E "passed|failed|error" gpurun_out/r02_gpu_suite_last.log | tail -2
timeout 300 python tools/cpu_vs_gpu_table.py > gpurun_out/r02_cpu_vs_gpu.md 2> gpurun_out/r02_cpu_vs_gpu.err; echo table rc=$?; cat gpurun_out/r02_cpu_vs_gpu.md
