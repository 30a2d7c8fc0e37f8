set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_prologue.py -m gpu -x -q > gpurun_out/pytest_prologue.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_prologue.log
tail -25 gpurun_out/pytest_prologue.log
python tools/probe_prologue.py --json gpurun_out/prologue_probe.json > gpurun_out/prologue_probe.log 2>&1
cat gpurun_out/prologue_probe.log
python tools/probe_prologue.py --n 131072 --hq 4 --hkv 1 --json gpurun_out/prologue_probe_p8.json > gpurun_out/prologue_probe_p8.log 2>&1
cat gpurun_out/prologue_probe_p8.log
python tools/probe_prologue.py --n 131072 --hq 8 --hkv 1 >> gpurun_out/prologue_probe_p8.log 2>&1
tail -3 gpurun_out/prologue_probe_p8.log
TMX_PROLOGUE_VARIANT=1 python tools/probe_prologue.py --n 131072 > gpurun_out/prologue_probe_v1.log 2>&1
TMX_PROLOGUE_VARIANT=1 python tools/probe_prologue.py --n 131072 --hq 4 --hkv 1 >> gpurun_out/prologue_probe_v1.log 2>&1
cat gpurun_out/prologue_probe_v1.log
ncu --set full --clock-control none --import-source on -k regex:prologue -s 4 -c 1 -o gpurun_out/prologue_128k python tools/probe_prologue.py --n 131072 > gpurun_out/ncu_full_prologue.log 2>&1
tail -3 gpurun_out/ncu_full_prologue.log
