set -x
mkdir -p gpurun_out
for rep in 1 2 3 4; do for v in 0 81; do
  echo "== TMX_SM100_VARIANT=$v"; TMX_SM100_VARIANT=$v timeout 300 python tools/probe.py --n 32768 131072 --kernels tcgen05 --modes 0 2>&1 | grep "^attn\|rror"
done; done > gpurun_out/probe_ab81.log 2>&1
cat gpurun_out/probe_ab81.log
TMX_SM100_VARIANT=81 timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_v81.log 2>&1; echo "variant 81 full suite rc=$?"; tail -4 gpurun_out/pytest_v81.log
