#!/bin/bash
# multi-GPU call (gpurun --gpus N): sharded-vs-single bit-identity test, then bench.py sharded and replicas at N ranks
N=${1:-2}
FILTER=${2:-}
mkdir -p gpurun_out
if [ -n "$FILTER" ]; then
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q -k "$FILTER" > gpurun_out/pytest_shard_$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_shard_$N.log
else
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q > gpurun_out/pytest_shard_$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_shard_$N.log
fi
tail -5 gpurun_out/pytest_shard_$N.log
run() { # name, args...
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 20 --warmup 5 "$@" 2> gpurun_out/bench_${name}_$N.err | grep '^{' > gpurun_out/bench_${name}_$N.json
  echo "$name exit $?"; cat gpurun_out/bench_${name}_$N.json; tail -3 gpurun_out/bench_${name}_$N.err
}
run sharded_1m --parallelism sharded --workload sc_1m
run sharded_4m --parallelism sharded --workload sc_4m
run replicas_1m --parallelism replicas --workload sc_1m
