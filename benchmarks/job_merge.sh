set -x
( time timeout 600 python -m pytest tests/test_gpu_listlist.py tests/test_cpp_facade.py tests/test_gpu_list.py -m gpu -q -x ) > gpurun_out/pytest_merge.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_merge.log
