set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
python tools/probe.py --n 32768 131072 --kernels auto --modes 1 0 > gpurun_out/probe.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_stream2 -s 2 -c 1 -o gpurun_out/stream2_tri_128k python tools/probe.py --n 131072 --kernels auto --modes 1 > gpurun_out/ncu_full_stream2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_sm100_v3 -s 2 -c 1 -o gpurun_out/tail_tc_128k python tools/probe.py --n 131072 --kernels auto --modes 1 > gpurun_out/ncu_full_tail.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/smoke.log; cat gpurun_out/bench.json; cat gpurun_out/probe.log
