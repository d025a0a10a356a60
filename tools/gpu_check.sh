#!/usr/bin/env bash
# One guarded GPU pass: every step runs under its own `timeout` and writes its log into gpurun_out/ as it goes, so a
# hung kernel costs one step's limit (not the whole call) and whatever finished before it is still pulled back.
#
#   gpurun --timeout 420 -- 'bash tools/gpu_check.sh quick'      # kernel tests of the step + a short bench
#   gpurun --timeout 900 -- 'bash tools/gpu_check.sh full'       # whole GPU suite, bench, reference arm, launch list
#   MGN_ALT_LIB=graph_physics_b200/libmgn_b200.exp.so bash tools/gpu_check.sh ab
#       `ab`: bench the built library, then swap in an experimental build of the library (kept next to it) and run
#       the step tests + the same bench on it -- a candidate that hangs cannot take the baseline numbers with it.
set -u
mode="${1:-quick}"
out=gpurun_out
mkdir -p "$out"
step() {   # step <name> <seconds> <command...>
  local name="$1" limit="$2"; shift 2
  echo "== $name (limit ${limit}s)" | tee -a "$out/check.log"
  timeout --signal=TERM --kill-after=10 "$limit" "$@" > "$out/$name.log" 2>&1
  local rc=$?
  echo "   rc=$rc  $(tail -n 1 "$out/$name.log" | cut -c1-300)" | tee -a "$out/check.log"
  return $rc
}
STEP_TESTS="tests/test_gpu_fwd_step.py tests/test_gpu_bwd_step.py tests/test_gpu_processor.py tests/test_gpu_encoder.py"
case "$mode" in
  quick)
    step tests_step 240 python -m pytest $STEP_TESTS -x -q
    step bench_short 120 python bench.py --steps 10 --warmup 3 --cpu-seconds 0
    ;;
  full)
    step tests_gpu 420 python -m pytest tests -m gpu -x -q
    step bench 240 python bench.py
    step bench_reference 240 python bench.py --impl reference
    step launches 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
         --log-file "$out/launches.csv" python bench.py --steps 2 --warmup 3 --cpu-seconds 0
    ;;
  ab)
    alt="${MGN_ALT_LIB:?set MGN_ALT_LIB to the experimental .so}"
    step bench_base 120 python bench.py --steps 10 --warmup 3 --cpu-seconds 0
    cp graph_physics_b200/libmgn_b200.so "$out/libmgn_b200.base.so.keep" && cp "$alt" graph_physics_b200/libmgn_b200.so
    step tests_alt 240 python -m pytest $STEP_TESTS -x -q && \
      step bench_alt 120 python bench.py --steps 10 --warmup 3 --cpu-seconds 0
    cp "$out/libmgn_b200.base.so.keep" graph_physics_b200/libmgn_b200.so; rm -f "$out/libmgn_b200.base.so.keep"
    ;;
  *) echo "usage: $0 quick|full|ab"; exit 2 ;;
esac
cat "$out/check.log"
