# instruction-cache counters of the streaming kernel (L1.5 = "GCC": requests are L0 misses)
mkdir -p gpurun_out
timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/plain3.log 2>&1 && \
timeout 600 ncu --clock-control none -k regex:plevel_stream -s 2 -c 1 --csv --log-file gpurun_out/icache.csv \
  --metrics gcc__cache_requests_type_instruction.sum,gcc__cache_requests_type_instruction_lookup_hit.sum,gcc__cache_requests_type_instruction_lookup_miss.sum,gcc__cache_requests_type_constant.sum,smsp__inst_executed.sum,smsp__inst_issued.sum,sm__cycles_elapsed.max,gpu__time_duration.sum,smsp__warp_issue_stalled_no_instruction_per_warp_active.pct \
  python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2 > gpurun_out/ncu_icache.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/icache.csv')) if r and not r[0].startswith('==')]
h=rows[0]
for r in rows[1:]:
    print(r[h.index('Metric Name')], r[h.index('Metric Value')], r[h.index('Metric Unit')])
PY
