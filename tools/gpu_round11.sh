set -x
mkdir -p gpurun_out
TMX_SM100_VARIANT=112 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "attention_vs_oracle or attention_golden or peaky or varlen" > gpurun_out/pytest_v112.log 2>&1; echo "variant 112 rc=$?"; tail -4 gpurun_out/pytest_v112.log
for v in 0 112 113 81 0 112 113; do
  echo "== TMX_SM100_VARIANT=$v"; TMX_SM100_VARIANT=$v timeout 300 python tools/probe.py --n 32768 131072 --kernels tcgen05 --modes 0 2>&1 | grep "^attn\|rror"
done > gpurun_out/probe_poll.log 2>&1
cat gpurun_out/probe_poll.log
