# Full GPU pass of a round: tests, smoke, bench (both arms), launch list and one full capture of the dominant kernel.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench_reference.json gpurun_out/bench.json
python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/ncu_launches.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:plevel_pair_kernel -s 2 -c 1 -o gpurun_out/prof_stream python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
# the other kernels of the path at full size: event timing from Python, then their ncu durations
python tools/kernel_rooflines.py > gpurun_out/kernel_rooflines.txt 2> gpurun_out/kernel_rooflines.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/kernels_ncu.csv python tools/kernel_rooflines.py > gpurun_out/kernels_ncu.log 2>&1
cat gpurun_out/kernel_rooflines.txt
# the stand-alone remapper: timing, then one full capture of the streaming kernel
python tools/mappm_bench.py 10 > gpurun_out/mappm_bench.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mappm_stream_kernel -s 3 -c 1 -o gpurun_out/prof_mappm python tools/mappm_bench.py 2 > gpurun_out/ncu_mappm.log 2>&1
cat gpurun_out/mappm_bench.log
