# timing experiments: rebuild the library with -DCG_EXP=n on the GPU box and time the bench
set -x
mkdir -p gpurun_out
for e in 5 6; do
  touch fv3net_b200/csrc/plevel_stream.cu
  make -C fv3net_b200/csrc NVFLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -DCG_EXP=$e" > gpurun_out/build_exp$e.log 2>&1
  python bench.py --steps 5 --warmup 3 --no-cpu --e2e-steps 0 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('EXP $e ms', d['roofline']['ms'])"
done
