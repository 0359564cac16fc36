set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01_a.json 2> gpurun_out/bench_r01_a.err; echo "bench exit $?"
./build/ubench_pipes > gpurun_out/ubench_pipes.txt 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/ncu_launches.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:plevel_fused -s 1 -c 1 -o gpurun_out/prof_fused python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_r01_a.json; cat gpurun_out/ubench_pipes.txt
