for prio in 0 1; do for rep in 1 2 3 4; do
  SDRX_MAIN_PRIO=$prio timeout 200 python bench.py --no-alt --no-cpu --sustain 0 --no-worst > gpurun_out/r.json 2> gpurun_out/r.err; rc=$?
  echo "main prio $prio rep $rep rc=$rc: $(python tools/bench_summary.py gpurun_out/r.json 2>/dev/null | head -1)"
done; done
for prio in 0 1; do SDRX_MAIN_PRIO=$prio timeout 100 python tools/diag_timeline.py --mode tc | tail -14; done
