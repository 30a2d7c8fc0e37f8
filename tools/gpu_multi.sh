# usage: bash tools/gpu_multi.sh P
set -x
P=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py --out gpurun_out/multi_gpu_check_n$P.json > gpurun_out/multi_gpu_check_n$P.log 2>&1; echo "check rc=$?"
grep -v "^W\|^\[W" gpurun_out/multi_gpu_check_n$P.log | tail -6
python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $P --steps 2 --warmup 3 > gpurun_out/bench_n$P.json 2> gpurun_out/bench_n$P.err; echo "bench rc=$?"
cat gpurun_out/bench_n$P.json; tail -3 gpurun_out/bench_n$P.err
