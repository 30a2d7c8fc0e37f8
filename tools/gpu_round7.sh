set -x
mkdir -p gpurun_out
for v in 0 64 65 66 0 64; do
  echo "== TMX_SM100_VARIANT=$v"; TMX_SM100_VARIANT=$v python tools/probe.py --n 32768 131072 --kernels tcgen05 --modes 0 2>&1 | grep "^attn"
done > gpurun_out/probe_pipe.log 2>&1
cat gpurun_out/probe_pipe.log
TMX_SM100_VARIANT=64 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "attention" > gpurun_out/pytest_pipe.log 2>&1; tail -3 gpurun_out/pytest_pipe.log
