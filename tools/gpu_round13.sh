set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
TMX_FUZZ_SEEDS=1000 python -m pytest tests/test_gpu_parity.py -m gpu -q -k fuzz > gpurun_out/pytest_fuzz1000.log 2>&1; tail -2 gpurun_out/pytest_fuzz1000.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
TMX_SM100_CLASSIC=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_classic.json 2> gpurun_out/bench_classic.err; echo "classic rc=$?"
python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
for f in ("bench_classic.json","bench.json"):
    d=json.load(open("gpurun_out/"+f)); print(f, d["value"], d["ms_per_layer"], d["roofline"]["frac"], d["clocks"]["sm_mhz"])
PY
ncu --set full --clock-control none --import-source on -k regex:attn_sm100_v3 -s 1 -c 1 -f -o gpurun_out/sm100_v3_default_128k python tools/probe.py --n 131072 --kernels auto --modes 0 > gpurun_out/ncu_full_default.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
tail -1 gpurun_out/ncu_full_default.log
