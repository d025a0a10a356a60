#!/usr/bin/env bash
# Liveness soak of the measured path (VERDICT r1 "next round" 1c): bench.py in FRESH processes, N runs per variant,
# each under its own timeout, one summary line per run in gpurun_out/soak.log.  bench.py's own watchdog and the bounded
# barrier waits of the kernels name the phase / the wait if a run stalls.
#   bash tools/soak.sh [runs_per_variant] [variants...]      variants: default pdl0 nograph full
set -u
runs="${1:-15}"; shift || true
variants=("$@"); [ ${#variants[@]} -eq 0 ] && variants=(default pdl0 nograph)
out=gpurun_out; mkdir -p "$out"
log="$out/soak.log"; : > "$log"
for v in "${variants[@]}"; do
  ok=0; bad=0
  for i in $(seq 1 "$runs"); do
    case "$v" in
      default) env=(); args=(--steps 20 --warmup 5 --cpu-seconds 0) ;;
      full)    env=(); args=(--steps 20 --warmup 5) ;;
      pdl0)    env=(MGN_PDL=0); args=(--steps 20 --warmup 5 --cpu-seconds 0) ;;
      nograph) env=(); args=(--steps 20 --warmup 5 --cpu-seconds 0 --no-cuda-graph) ;;
      *) echo "unknown variant $v"; exit 2 ;;
    esac
    t0=$(date +%s.%N)
    env "${env[@]}" timeout --signal=TERM --kill-after=10 150 python bench.py "${args[@]}" --phase-timeout 60 \
        > "$out/soak_${v}_$i.out" 2> "$out/soak_${v}_$i.err"
    rc=$?
    t1=$(date +%s.%N)
    ms=$(python - "$out/soak_${v}_$i.out" <<'PY'
import json,sys
try:
    l=[x for x in open(sys.argv[1]).read().splitlines() if x.startswith("{")][-1]
    d=json.loads(l); print("%.3f"%d.get("ms_per_step",-1), d.get("error",""))
except Exception as e:
    print("nan", "no line")
PY
)
    printf "%s run %2d rc=%d wall=%.1fs ms_per_step=%s\n" "$v" "$i" "$rc" "$(echo "$t1 - $t0" | bc)" "$ms" | tee -a "$log"
    if [ "$rc" -eq 0 ]; then ok=$((ok+1)); rm -f "$out/soak_${v}_$i.err" "$out/soak_${v}_$i.out"; else bad=$((bad+1)); tail -n 5 "$out/soak_${v}_$i.err" | tee -a "$log"; fi
  done
  echo "== $v: $ok ok, $bad failed of $runs" | tee -a "$log"
done
