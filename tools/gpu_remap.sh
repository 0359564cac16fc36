mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_remap.py -m gpu -q --tb=short -x > gpurun_out/pytest_plevel.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_plevel.log
tail -30 gpurun_out/pytest_plevel.log
