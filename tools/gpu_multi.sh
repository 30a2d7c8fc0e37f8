#!/bin/bash
# usage (under gpurun --gpus P):  bash tools/gpu_multi.sh P <stage> [<stage> ...]
#   check[:N]       multi-GPU parity of the head-sharded path against the 1-GPU result (tools/multi_gpu_check.py)
#   bench:<config>  bench.py --config <2|3|4> on P GPUs (torchrun), JSON line into gpurun_out/bench_config<c>_n<P>.json
P=$1; shift
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
for stage in "$@"; do
  name=${stage%%:*}; arg=""; [[ "$stage" == *:* ]] && arg=${stage#*:}
  echo "=== P=$P stage $name $arg"
  case $name in
    check)
      timeout 900 $RUN --master-port 29511 tools/multi_gpu_check.py --tokens ${arg:-131072} --out gpurun_out/multi_gpu_check_n$P.json \
        > gpurun_out/multi_gpu_check_n$P.log 2>&1; echo "check rc=$?"
      grep -v "^W\|^\[W" gpurun_out/multi_gpu_check_n$P.log | tail -6 ;;
    bench)
      c=${arg:-2}
      timeout 900 $RUN --master-port 29512 bench.py --gpus $P --config $c --steps ${STEPS:-3} --warmup 3 \
        > gpurun_out/bench_config${c}_n$P.json 2> gpurun_out/bench_config${c}_n$P.err; echo "bench rc=$?"
      cat gpurun_out/bench_config${c}_n$P.json; grep -v "^W\|^\[W" gpurun_out/bench_config${c}_n$P.err | tail -3 ;;
  esac
done
