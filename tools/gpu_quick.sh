set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python tools/probe.py --n 32768 131072 --kernels auto --modes 1 > gpurun_out/probe.log 2>&1
tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/probe.log
