mkdir -p gpurun_out
timeout 900 python -X faulthandler -m pytest tests/test_gpu_staging.py -x -q -s > gpurun_out/r02_staging_test.log 2>&1; echo rc=$?
grep -n "passed\|failed\|Fatal\|Error\|H2D" gpurun_out/r02_staging_test.log | head -20
grep -n "File \"/root\|File \"/tmp\|tests/" gpurun_out/r02_staging_test.log | head -20
