#!/bin/bash
# One parameterised GPU lease script:  gpurun -- 'bash tools/gpu_run.sh <stage> [<stage> ...]'
# Every stage logs into gpurun_out/ (merged back by gpurun) and prints a short tail.
# Stages: tests | tests:<pytest -k expr> | bench | bench:<extra args> | probe:<probe.py args> | tri[:N] |
#         ncu_tri[:N] | ncu_dense[:N] | sanitize | compare | launches
mkdir -p gpurun_out
T=${TMX_STAGE_TIMEOUT:-900}
for stage in "$@"; do
  name=${stage%%:*}; arg=""; [[ "$stage" == *:* ]] && arg=${stage#*:}
  echo "=== stage $name $arg"
  case $name in
    tests)
      timeout $T python -m pytest tests -m gpu -x -q ${arg:+-k "$arg"} > gpurun_out/pytest_gpu.log 2>&1
      echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log ;;
    smoke)
      timeout $T python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -8 gpurun_out/smoke.log ;;
    bench)
      tag=$(echo "$arg" | tr -c 'a-zA-Z0-9\n' '_')
      timeout $T python bench.py $arg > gpurun_out/bench$tag.json 2> gpurun_out/bench$tag.err; echo "bench rc=$?"
      cat gpurun_out/bench$tag.json; tail -3 gpurun_out/bench$tag.err ;;
    probe)
      timeout $T python tools/probe.py $arg 2>&1 | tee -a gpurun_out/probe.log ;;
    tri)
      # per-kernel times of a triangle layer: CUDA events, then the ncu launch list (cold-cache, serialised)
      N=${arg:-131072}
      timeout $T python tools/probe.py --n $N --kernels auto --modes 1 2>&1 | grep attn | tee -a gpurun_out/probe.log
      timeout $T ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tri_launches.csv \
        python tools/probe.py --n $N --kernels auto --modes 1 > /dev/null 2>&1
      python tools/ncu_summary.py --launches gpurun_out/tri_launches.csv | tee gpurun_out/tri_launches.txt ;;
    ncu_tri)
      N=${arg:-131072}
      timeout $T ncu --set full --clock-control none --import-source on -k regex:attn_stream3 -s 2 -c 1 -f \
        -o gpurun_out/stream3_tri python tools/probe.py --n $N --kernels auto --modes 1 > gpurun_out/ncu_tri.log 2>&1
      echo "ncu rc=$?"; tail -3 gpurun_out/ncu_tri.log ;;
    ncu_dense)
      N=${arg:-131072}
      timeout $T ncu --set full --clock-control none --import-source on -k regex:attn_sm100 -s 1 -c 1 -f \
        -o gpurun_out/dense python tools/probe.py --n $N --kernels auto --modes 0 > gpurun_out/ncu_dense.log 2>&1
      echo "ncu rc=$?"; tail -3 gpurun_out/ncu_dense.log ;;
    sanitize)
      # memcheck + racecheck + synccheck on small shapes of every kernel family
      for tool in memcheck racecheck synccheck; do
        timeout $T compute-sanitizer --tool $tool --error-exitcode 9 python tools/sanitize_cases.py \
          > gpurun_out/sanitizer_$tool.log 2>&1
        echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|cases ok" gpurun_out/sanitizer_$tool.log | tail -3
      done ;;
    compare)
      timeout $T python tools/compare_dense.py > gpurun_out/compare_dense.log 2>&1; echo "compare rc=$?"; tail -12 gpurun_out/compare_dense.log ;;
    launches)
      timeout $T ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
        python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "launches rc=$?"
      python tools/ncu_summary.py --launches gpurun_out/launches.csv | tee gpurun_out/launches.txt ;;
    *) echo "unknown stage $name" ;;
  esac
done
