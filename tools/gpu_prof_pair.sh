# one full ncu capture of the pair-thread kernel (after a plain run of the same command exits 0)
mkdir -p gpurun_out
timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:plevel_pair_kernel -s 2 -c 1 -f -o gpurun_out/prof_pair python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
