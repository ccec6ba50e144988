mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
SAN=/usr/local/cuda/bin/compute-sanitizer
( time timeout 1200 $SAN --tool memcheck --target-processes all --error-exitcode 7 --print-limit 20 --log-file gpurun_out/sanitizer_memcheck.log \
  python -m pytest tests/test_gpu_census_sgm.py tests/test_gpu_float_cv.py tests/test_gpu_refine.py tests/test_gpu_filters.py tests/test_gpu_image_refinement.py tests/test_zz_no_alloc.py -m gpu -x -q -k "not fullsize" ) > gpurun_out/sanitizer_memcheck_pytest.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck_pytest.log
tail -5 gpurun_out/sanitizer_memcheck_pytest.log; tail -5 gpurun_out/sanitizer_memcheck.log
( time timeout 1200 $SAN --tool racecheck --target-processes all --error-exitcode 7 --print-limit 20 --log-file gpurun_out/sanitizer_racecheck.log \
  python -m pytest tests/test_gpu_float_cv.py tests/test_gpu_sgm_general.py tests/test_gpu_filters.py -m gpu -x -q -k "not fullsize" ) > gpurun_out/sanitizer_racecheck_pytest.log 2>&1
echo "racecheck rc=$?" >> gpurun_out/sanitizer_racecheck_pytest.log
tail -5 gpurun_out/sanitizer_racecheck_pytest.log; tail -5 gpurun_out/sanitizer_racecheck.log
