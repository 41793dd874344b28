python tools/cpu_vs_gpu_table.py > gpurun_out/cpu_vs_gpu.md 2> gpurun_out/cpu_vs_gpu.err; cat gpurun_out/cpu_vs_gpu.md; tail -3 gpurun_out/cpu_vs_gpu.err
