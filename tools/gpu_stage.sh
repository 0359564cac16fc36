mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_staging.py -m gpu -q --tb=short > gpurun_out/pytest_staging.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_staging.log
tail -30 gpurun_out/pytest_staging.log
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 5 > gpurun_out/bench_stage.json 2> gpurun_out/bench_stage.err; echo "bench exit $?"; tail -3 gpurun_out/bench_stage.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_stage.json').read().strip().splitlines()[-1])
print("ms_per_step", d["ms_per_step"], "plevel ms", d["roofline"]["ms"], "e2e", d["e2e"])
PY
