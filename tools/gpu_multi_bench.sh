set -x
P=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $P --steps 2 --warmup 3 > gpurun_out/bench_n$P.json 2> gpurun_out/bench_n$P.err; echo "bench rc=$?"
echo "stdout lines: $(wc -l < gpurun_out/bench_n$P.json)"; cat gpurun_out/bench_n$P.json; tail -3 gpurun_out/bench_n$P.err
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench1 rc=$?"; cut -c1-400 gpurun_out/bench_quick.json
