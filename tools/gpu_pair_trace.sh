# timeline of one CTA: rebuild with the trace points (+ flags given as arguments, one variant each), run, summarise
mkdir -p gpurun_out
i=0
for e in "$@"; do
  i=$((i+1))
  touch fv3net_b200/csrc/plevel_pair.cu
  make -C fv3net_b200/csrc EXTRA="-DCG_PAIR_TRACE $e" > gpurun_out/build_trace.log 2>&1 || { tail -5 gpurun_out/build_trace.log; exit 1; }
  timeout 300 python tools/pair_trace.py > gpurun_out/pair_trace$i.txt 2>&1; echo "variant [$e] exit $?"
  python tools/pair_trace_summary.py gpurun_out/pair_trace$i.txt
done
