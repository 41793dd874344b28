# the no-allocation entry point: its test alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_staging.py -x -q -s 2>&1 | tail -8
python tools/first_call.py 2>&1 | grep -E "default|pin 128" | tee gpurun_out/r02_first_call_after.txt
( time timeout 1700 python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_gpu_suite_g1.log 2>&1; tail -6 gpurun_out/r02_gpu_suite_g1.log
