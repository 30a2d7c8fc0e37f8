set -x
mkdir -p gpurun_out
TMX_SM100_VARIANT=80 ncu --set full --clock-control none --import-source on -k regex:attn_sm100_v3 -s 1 -c 1 -o gpurun_out/sm100_v3_ahead_128k python tools/probe.py --n 131072 --kernels tcgen05 --modes 0 > gpurun_out/ncu_full_ahead.log 2>&1
tail -2 gpurun_out/ncu_full_ahead.log
