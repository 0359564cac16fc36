# quick check of the pressure-level path: parity tests, then a short device-resident bench
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_remap.py -m gpu -q -x > gpurun_out/pytest_plevel.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_plevel.log
tail -8 gpurun_out/pytest_plevel.log
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/bench_stream.json 2> gpurun_out/bench_stream.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_stream.json').read().strip().splitlines()[-1])
print("ms_per_step", d["ms_per_step"], "plevel ms", d["roofline"]["ms"], "frac", d["roofline"]["frac"])
PY
