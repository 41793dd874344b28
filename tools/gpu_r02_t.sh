mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_multi.py -x -q ) > gpurun_out/r02_multi_test.log 2>&1
tail -60 gpurun_out/r02_multi_test.log
