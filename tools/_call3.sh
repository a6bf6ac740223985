mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2c3_noalloc_alone.log 2>&1; echo "rc $?" >> gpurun_out/r2c3_noalloc_alone.log
tail -4 gpurun_out/r2c3_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2c3_tests.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2c3_tests.log
tail -4 gpurun_out/r2c3_tests.log
