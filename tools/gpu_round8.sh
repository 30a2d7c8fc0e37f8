set -x
mkdir -p gpurun_out
TMX_SM100_VARIANT=80 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "attention_vs_oracle or attention_golden" > gpurun_out/pytest_ahead.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_ahead.log
for v in 0 80 0 80; do
  echo "== TMX_SM100_VARIANT=$v"; TMX_SM100_VARIANT=$v timeout 300 python tools/probe.py --n 32768 131072 --kernels tcgen05 --modes 0 2>&1 | grep "^attn\|rror"
done > gpurun_out/probe_ahead.log 2>&1
cat gpurun_out/probe_ahead.log
