#!/bin/bash
# One parameterised GPU-box runner (replaces the round-1 run_*.sh scratch launchers).
#   tools/run.sh <tag> <steps...>    steps: noalloc tests smoke bench benchref launches ncufull case:<name>:<batch>
mkdir -p gpurun_out
T=${1:-r2}; shift
for step in "$@"; do
case $step in
noalloc)
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/${T}_noalloc_alone.log 2>&1
  echo "noalloc alone exit $?"; tail -n 3 gpurun_out/${T}_noalloc_alone.log ;;
tests)
  timeout 1800 python -m pytest tests -x -q -m gpu --timeout 400 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
  echo "tests exit $?" | tee -a gpurun_out/${T}_tests.log; grep -v -i warning gpurun_out/${T}_tests.log | tail -n 6 ;;
smoke)
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${T}_smoke.log ;;
bench)
  timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"
  python tools/show_bench.py gpurun_out/${T}_bench.json ;;
benchref)
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"
  tail -n 1 gpurun_out/${T}_bench_ref.json | cut -c1-400 ;;
launches)
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${T}_ncu.log 2>&1
  echo "ncu exit $?" ;;
ncufull)
  # one DeiT-S block of the second forward (patch GEMM, LN, qkv, attention, proj, LN, fc1, GELU, fc2), full sections
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:"gemm_i8_|layernorm_|attention_i8|gelu_lut_apply" -s 99 -c 9 \
    -o gpurun_out/${T}_block -f python tools/prof_step.py > gpurun_out/${T}_ncufull.log 2>&1
  echo "ncu full exit $?" ;;
others)
  # the other BASELINE.json configs on the current build -> one JSON list
  echo "[" > gpurun_out/${T}_bench_other_configs.json
  first=1
  for cb in "deit_tiny_ivit 8" "deit_tiny_ivit 256" "vit_base_mixed 64" "vit_base_mixed 256" "swin_tiny 256" "deit_base_ibert_w16 64" "deit_base_ibert_w16 256" "deit_base_ibert 64" "deit_tiny_ppoly 256"; do
    set -- $cb
    timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --case $1 --batch $2 > gpurun_out/_o.json 2> gpurun_out/_o.err || { echo "bench $1 $2 failed"; tail -n 3 gpurun_out/_o.err; continue; }
    [ $first = 1 ] || echo "," >> gpurun_out/${T}_bench_other_configs.json
    first=0
    tail -n 1 gpurun_out/_o.json >> gpurun_out/${T}_bench_other_configs.json
    python - <<PY
import json
d=json.loads(open("gpurun_out/_o.json").read().strip().splitlines()[-1])
print("$1 $2", "value", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
PY
  done
  echo "]" >> gpurun_out/${T}_bench_other_configs.json ;;
case:*)
  IFS=: read -r _ name batch <<< "$step"
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --case $name --batch $batch > gpurun_out/${T}_bench_$name.json 2> gpurun_out/${T}_bench_$name.err
  echo "bench $name $batch exit $?"; python tools/show_bench.py gpurun_out/${T}_bench_$name.json ;;
*)
  echo "running: $step"; bash -c "$step" ;;
esac
done
