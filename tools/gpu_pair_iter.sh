# one iteration on the pair-thread kernel: its parity tests, then a short device-resident bench
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_remap.py -m gpu -q --tb=line -x -k "pair or default" > gpurun_out/pytest_pair.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_pair.log
tail -8 gpurun_out/pytest_pair.log
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench exit $?"
tail -3 gpurun_out/bench_iter.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_iter.json').read().strip().splitlines()[-1])
print("ms_per_step", d["ms_per_step"], "plevel ms", d["roofline"]["ms"], "frac", d["roofline"]["frac"])
PY
