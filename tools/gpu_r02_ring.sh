# round 2: small staging rings cut into fewer, larger slots -- the GPU suite, then first / later upload per size
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gpu_suite_ring.log 2>&1; grep -E "passed|failed|error" gpurun_out/r02_gpu_suite_ring.log | tail -2
timeout 300 python tools/first_call.py 4096 8192 12288 16384 > gpurun_out/r02_first_call_slots.txt 2>&1; echo rc=$?; grep -v "CHUNK_KB\|STAGE_THREADS" gpurun_out/r02_first_call_slots.txt
