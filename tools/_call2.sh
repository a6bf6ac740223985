set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2c2_noalloc_alone.log 2>&1; echo "rc $?" >> gpurun_out/r2c2_noalloc_alone.log
tail -15 gpurun_out/r2c2_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2c2_tests.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2c2_tests.log
tail -15 gpurun_out/r2c2_tests.log
timeout 600 tools/sanitize.sh r2c2 memcheck -- lin_dare general > gpurun_out/r2c2_sanitize_stdout.txt 2>&1
tail -30 gpurun_out/r2c2_sanitize_stdout.txt; cat gpurun_out/sanitizer_r2c2_memcheck.txt | head -60
