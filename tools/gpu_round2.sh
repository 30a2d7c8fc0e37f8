set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
ncu --set full --clock-control none --import-source on -k regex:attn_sm100_v3 -s 1 -c 1 -o gpurun_out/sm100_v3_dense_128k python tools/probe.py --n 131072 --kernels auto --modes 0 > gpurun_out/ncu_full_v3_128k.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_stream -s 2 -c 1 -o gpurun_out/stream_tri_128k python tools/probe.py --n 131072 --kernels auto --modes 1 > gpurun_out/ncu_full_stream.log 2>&1
ncu --set full --clock-control none -k regex:mask_bits -s 1 -c 1 -o gpurun_out/mask_bits_128k python tools/probe.py --n 131072 --kernels auto --modes 1 > gpurun_out/ncu_full_mask.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
cat gpurun_out/bench.json
