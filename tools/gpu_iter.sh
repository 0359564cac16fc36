# one iteration on the streaming kernel: its parity tests (+ full-size tests), then a short device-resident bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_remap.py tests/test_gpu_fullsize.py -m gpu -q -x --tb=short > gpurun_out/pytest_iter.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_iter.log
tail -25 gpurun_out/pytest_iter.log
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_iter.json').read().strip().splitlines()[-1])
print("ms_per_step", d["ms_per_step"], "plevel ms", d["roofline"]["ms"], "frac", d["roofline"]["frac"])
PY
