set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_remap.py -m gpu -q -k "pressure_level" > gpurun_out/pytest_plevel.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_plevel.log
tail -40 gpurun_out/pytest_plevel.log
