# round 2, text ingest after the SSE2 token locator: the GPU suite, then the ingest rate with the library's phase stamps
mkdir -p gpurun_out
nproc
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gpu_suite_ingest.log 2>&1; tail -3 gpurun_out/r02_gpu_suite_ingest.log
timeout 300 python tools/ingest_rate.py 8192 8 32 4 16 2 8 > gpurun_out/r02_ingest_n8192.json 2> gpurun_out/r02_ingest_n8192.err; echo ingest rc=$?
cat gpurun_out/r02_ingest_n8192.json; grep -c . gpurun_out/r02_ingest_n8192.err; awk 'NR%3==0' gpurun_out/r02_ingest_n8192.err | head -8
timeout 300 python tools/ingest_rate.py 16384 8 32 > gpurun_out/r02_ingest_n16384.json 2> gpurun_out/r02_ingest_n16384.err; echo ingest rc=$?
cat gpurun_out/r02_ingest_n16384.json; tail -2 gpurun_out/r02_ingest_n16384.err
