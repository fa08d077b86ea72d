#!/usr/bin/env bash
# Multi-GPU checks on N = $1 GPUs of one box: KV-transfer tests, the C5 harness, bench.py weak scaling.
set -uo pipefail
N=${1:-2}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo_$N.txt 2>&1
python -m pytest tests/test_kv_transfer_gpu.py -x -q > $OUT/pytest_kv_$N.log 2>&1; echo "pytest kv exit $?"; tail -5 $OUT/pytest_kv_$N.log
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $RUN tools/kv_transfer_bench.py --prompts ${PROMPTS:-8} > $OUT/kv_transfer_$N.jsonl 2> $OUT/kv_transfer_$N.err
echo "kv bench exit $?"; cat $OUT/kv_transfer_$N.jsonl | cut -c1-330; tail -5 $OUT/kv_transfer_$N.err
for w in ${WORKLOADS:-c2 c3tp8}; do
  timeout 600 $RUN bench.py --gpus $N --workload $w --steps 200 --warmup 5 > $OUT/bench_${w}_n$N.json 2> $OUT/bench_${w}_n$N.err
  echo "bench $w N=$N exit $?"; cut -c1-400 $OUT/bench_${w}_n$N.json; tail -3 $OUT/bench_${w}_n$N.err
done
