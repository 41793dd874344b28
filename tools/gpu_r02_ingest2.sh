# round 2, text ingest with the persistent copy team: ingest tests, rate per chunk size, per-chunk threads for comparison
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_ingest.py tests/test_gpu_reference_callers.py -x -q ) > gpurun_out/r02_ingest_tests.log 2>&1; tail -2 gpurun_out/r02_ingest_tests.log
timeout 300 python tools/ingest_rate.py 8192 8 4 2 16 8 4 > gpurun_out/r02_ingest_team_n8192.json 2> gpurun_out/r02_ingest_team_n8192.err; echo rc=$?
cat gpurun_out/r02_ingest_team_n8192.json; awk 'NR%3==0' gpurun_out/r02_ingest_team_n8192.err | head -8
CGB_INGEST_NO_POOL=1 timeout 300 python tools/ingest_rate.py 8192 8 4 > gpurun_out/r02_ingest_nopool_n8192.json 2> gpurun_out/r02_ingest_nopool_n8192.err; echo rc=$?
cat gpurun_out/r02_ingest_nopool_n8192.json; awk 'NR%3==0' gpurun_out/r02_ingest_nopool_n8192.err | head -3
timeout 300 python tools/ingest_rate.py 16384 8 4 16 > gpurun_out/r02_ingest_team_n16384.json 2> gpurun_out/r02_ingest_team_n16384.err; echo rc=$?
cat gpurun_out/r02_ingest_team_n16384.json; awk 'NR%3==0' gpurun_out/r02_ingest_team_n16384.err | head -4
