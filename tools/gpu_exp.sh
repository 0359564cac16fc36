# timing experiments: rebuild the streaming kernel with extra -D flags on the GPU box and time the device-resident bench
# usage: bash tools/gpu_exp.sh "<flags of variant 1>" "<flags of variant 2>" ...   ("" = as committed)
mkdir -p gpurun_out
i=0
for e in "$@"; do
  i=$((i+1))
  touch fv3net_b200/csrc/plevel_stream.cu fv3net_b200/csrc/plevel_pair.cu
  make -C fv3net_b200/csrc EXTRA="$e" > gpurun_out/build_exp$i.log 2>&1 || { echo "EXP [$e] build failed"; tail -5 gpurun_out/build_exp$i.log; continue; }
  timeout 120 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 --config 2 2> gpurun_out/exp$i.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('EXP [$e] plevel ms', round(d['roofline']['ms'],3), 'step ms', round(d['ms_per_step'],3))"
done
