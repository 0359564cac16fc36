set -x
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_stream.json 2> gpurun_out/bench_stream.err; echo "bench exit $?"
cat gpurun_out/bench_stream.json; tail -5 gpurun_out/bench_stream.err
python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:plevel_stream -s 1 -c 1 -o gpurun_out/prof_stream python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/ncu_stream.log 2>&1
tail -3 gpurun_out/ncu_stream.log
