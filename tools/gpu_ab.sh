# A/B: time the bench with the library as built here, then with the library rebuilt from git HEAD's plevel_pair.cu
mkdir -p gpurun_out
run() { timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu --e2e-steps 0 --config 2 2> gpurun_out/ab.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 plevel ms', round(d['roofline']['ms'],3), 'step ms', round(d['ms_per_step'],3))"; }
run work; run work
cp fv3net_b200/csrc/plevel_pair.cu /tmp/work_pair.cu
cp tools/head_pair.cu fv3net_b200/csrc/plevel_pair.cu
make -C fv3net_b200/csrc > gpurun_out/build_ab.log 2>&1 || tail -5 gpurun_out/build_ab.log
run head; run head
cp /tmp/work_pair.cu fv3net_b200/csrc/plevel_pair.cu
