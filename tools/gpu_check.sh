#!/usr/bin/env bash
# One gpurun call: GPU parity tests, bench lines for every single-GPU workload, then the two ncu
# passes of the profiling recipe (launch list + one full capture of the attention kernel).
#   gpurun --timeout 2400 -- 'bash tools/gpu_check.sh'
# Everything lands in gpurun_out/ (scratch); summaries worth keeping are copied to profiles/ by hand.
set -uo pipefail
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/gpu.txt 2>&1

if [ "${SKIP_TESTS:-0}" != "1" ]; then
  python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1
  echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
  tail -5 $OUT/pytest_gpu.log
fi

python bench.py > $OUT/bench_c2_v1.json 2> $OUT/bench_c2_v1.err; echo "bench c2 v1 exit $?"; cat $OUT/bench_c2_v1.json
for wl in c2:v2 c3tp8:v1 c3tp4:v1 c3tp2:v1 c3:v1 c4:v1 c4:v2 c1:v1 refbench:v1 refbench:v2; do
  w=${wl%%:*}; v=${wl##*:}
  python bench.py --workload $w --version $v --steps 200 --warmup 5 --no-cpu-baseline \
      > $OUT/bench_${w}_${v}.json 2> $OUT/bench_${w}_${v}.err
  echo "bench $w $v exit $?"; cat $OUT/bench_${w}_${v}.json
done

if [ "${SKIP_NCU:-0}" != "1" ]; then
  CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
  $CMD > $OUT/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches.csv $CMD \
      > $OUT/ncu_launches.log 2>&1
  echo "ncu launches exit $?"
  $CMD > $OUT/plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:paged_attn_kernel -s 4 -c 2 -f -o $OUT/prof_attn_c2 $CMD \
      > $OUT/ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
