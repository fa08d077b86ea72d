#!/usr/bin/env bash
# r05 (round 2) first GPU call: the whole GPU suite, smoke(), the driver's bench command (with the `extra`
# blocks), the cache-op lines, one ncu --set full capture per cache kernel, and compute-sanitizer memcheck
# on the decode kernels.
#   gpurun --timeout 1500 -- 'bash tools/gpu_r05_verify.sh'
set -uo pipefail
OUT=gpurun_out/r05
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/gpu.txt 2>&1

timeout 600 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
tail -15 $OUT/pytest_gpu.log

timeout 120 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -3 $OUT/smoke.log

timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_driver_cmd.json 2> $OUT/bench_driver_cmd.err
echo "bench exit $?"; cat $OUT/bench_driver_cmd.json; tail -5 $OUT/bench_driver_cmd.err

timeout 300 python tools/cache_ops_bench.py --iters 30 > $OUT/cache_ops.jsonl 2> $OUT/cache_ops.err
echo "cache_ops exit $?"; tail -3 $OUT/cache_ops.err

# ncu full on the cache kernels (one launch each) -- verdict item 7 asks for dram bytes / algorithmic bytes
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'reshape_and_cache_tile|copy_blocks_kernel|to_fp8|from_fp8' \
    -c 40 -f -o $OUT/prof_cache_ops python tools/cache_ops_bench.py --iters 1 --ncu --only reshape,copy,swap,convert > $OUT/ncu_cache.log 2>&1
echo "ncu cache exit $?"; tail -3 $OUT/ncu_cache.log

# compute-sanitizer memcheck on the decode kernels (SIMT, MMA, split v1, v2), small shapes
for spec in "c2 v1 4 1024" "c3tp8 v1 8 1024" "c4 v1 4 2048" "c2 v2 4 1024" "c4 v2 4 2048" "refbench v1 2 4096"; do
  set -- $spec
  timeout 300 compute-sanitizer --tool memcheck --print-limit 5 python tools/attn_once.py --workload $1 --version $2 --S $3 --ctx $4 \
      --iters 1 --rc --check > $OUT/sanitizer_memcheck_$1_$2.txt 2>&1
  echo "memcheck $1 $2 exit $?"; tail -2 $OUT/sanitizer_memcheck_$1_$2.txt
done
