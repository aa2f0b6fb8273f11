#!/bin/bash
# multi-GPU check of a sharding change: bit-identity tests, then the sharded sc_1m bench line at N ranks
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q > gpurun_out/pytest_shard_$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_shard_$N.log
tail -4 gpurun_out/pytest_shard_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus $N --steps 20 --warmup 5 --parallelism sharded --workload sc_1m --configs none --no-cpu-baseline 2> gpurun_out/bench_sharded_1m_$N.err | grep '^{' > gpurun_out/bench_sharded_1m_$N.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_sharded_1m_$N.json'))
print('value %.3e'%d['value'], 'ms %.3f'%d['ms_per_step'], d['breakdown_ms_per_step'], 'e2e %.3e'%d['e2e']['value'])
PY
tail -2 gpurun_out/bench_sharded_1m_$N.err
