set -x
mkdir -p gpurun_out
for v in 80 81 82 83 65 66; do
  echo "== TMX_SM100_VARIANT=$v"; TMX_SM100_VARIANT=$v timeout 300 python tools/probe.py --n 32768 131072 --kernels tcgen05 --modes 0 2>&1 | grep "^attn\|rror"
done > gpurun_out/probe_ahead2.log 2>&1
cat gpurun_out/probe_ahead2.log
TMX_SM100_VARIANT=80 ncu --set full --clock-control none --import-source on -k regex:attn_sm100_v3 -s 1 -c 1 -f -o gpurun_out/sm100_v3_ahead_128k python tools/probe.py --n 131072 --kernels tcgen05 --modes 0 > gpurun_out/ncu_full_ahead.log 2>&1
tail -2 gpurun_out/ncu_full_ahead.log
