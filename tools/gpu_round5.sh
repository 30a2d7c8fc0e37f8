set -x
mkdir -p gpurun_out
timeout 1200 python tools/compare_dense.py --json gpurun_out/compare_dense.json > gpurun_out/compare_dense.log 2>&1; echo "compare rc=$?"
grep "^{" gpurun_out/compare_dense.log
tail -5 gpurun_out/compare_dense.log
ncu --set full --clock-control none --import-source on -k regex:prologue -s 4 -c 1 -o gpurun_out/prologue_128k python tools/probe_prologue.py --n 131072 > gpurun_out/ncu_full_prologue.log 2>&1
tail -2 gpurun_out/ncu_full_prologue.log
