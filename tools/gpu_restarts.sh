# restart recipes against the golden JSONs (no -x: show every failing configuration)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_restarts.py -m gpu -q --tb=short > gpurun_out/pytest_restarts.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_restarts.log
tail -120 gpurun_out/pytest_restarts.log
