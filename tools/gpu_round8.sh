set -x
mkdir -p gpurun_out
for v in 64 80; do
TMX_SM100_VARIANT=$v timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "attention_vs_oracle or attention_golden or peaky or varlen" > gpurun_out/pytest_v$v.log 2>&1; echo "variant $v rc=$?"; tail -4 gpurun_out/pytest_v$v.log
done
for v in 0 64 80 0 64 80; do
  echo "== TMX_SM100_VARIANT=$v"; TMX_SM100_VARIANT=$v timeout 300 python tools/probe.py --n 32768 131072 --kernels tcgen05 --modes 0 2>&1 | grep "^attn\|rror"
done > gpurun_out/probe_ahead.log 2>&1
cat gpurun_out/probe_ahead.log
