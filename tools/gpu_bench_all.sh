# the default bench line with every sub-record, summarised
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 "$@" > gpurun_out/bench_all.json 2> gpurun_out/bench_all.err; echo exit $?
tail -5 gpurun_out/bench_all.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_all.json").read().strip().splitlines()[-1])
print("configs[2] ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "impl", d["roofline"]["impl"])
print("e2e", d["e2e"])
for k, v in d.get("configs", {}).items():
    print(k, round(v["ms_per_step"], 3), "ms", "%.3g" % v["value"], "cells/s frac", round(v["roofline"]["frac"], 3),
          {kk: vv for kk, vv in v.items() if kk in ("kernel_ms", "collective", "shards_per_rank")})
print("strong", d.get("strong"))
PY
