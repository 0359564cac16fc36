# the whole GPU suite, quietly
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -v "^\.\+\s*\[" gpurun_out/pytest_gpu.log | tail -40
