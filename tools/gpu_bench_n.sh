# bench.py under torchrun on N GPUs of the box: bash tools/gpu_bench_n.sh N [bench flags]
N=$1; shift
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo exit $?
tail -5 gpurun_out/bench_n$N.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
print("configs[2] ms", d["ms_per_step"], "value %.4g" % d["value"], "frac", d["roofline"]["frac"])
print("e2e", d["e2e"])
for k, v in d.get("configs", {}).items():
    print(k, round(v["ms_per_step"], 3), "ms", "%.4g" % v["value"], "cells/s frac", round(v["roofline"]["frac"], 3),
          {kk: vv for kk, vv in v.items() if kk in ("kernel_ms", "collective", "shards_per_rank")})
s = d.get("strong")
if s: print("strong", round(s["ms_per_step"], 3), "ms value %.4g" % s["value"], "bit", s["bit_identical_to_unsharded"], s["shards_per_rank"], "kernel_ms", s["kernel_ms"], s["collective"], "frac", s["roofline"]["frac"])
PY
