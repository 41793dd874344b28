mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_reference_callers.py tests/test_gpu_ingest.py -x -q ) > gpurun_out/r02_multi_test.log 2>&1
tail -40 gpurun_out/r02_multi_test.log
